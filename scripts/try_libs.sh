#!/bin/bash
# usage: scripts/try_libs.sh lib1.so lib2.so ...   -> times variant 1 of each library on the bench workload
for l in "$@"; do echo "== $l"; LBGCMB_LIB=$PWD/$l python scripts/n0_variants.py 1 2>&1 | grep kernel_ms; done
