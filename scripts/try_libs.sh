#!/bin/bash
# usage: scripts/try_libs.sh lib1.so lib2.so ...   -> times the default N0 variant of each library on the bench workload
for l in "$@"; do echo "== $l"; LBGCMB_LIB=$PWD/$l python scripts/n0_variants.py ${N0V:-3} 2>&1 | grep kernel_ms; done
