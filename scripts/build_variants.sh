#!/bin/bash
# Build liblbgcmb variants for kernel experiments: scripts/build_variants.sh "name:flags" ...  -> build/lib_<name>.so
set -e
cd "$(dirname "$0")/../lbgcmb_b200/csrc"
mkdir -p ../../build
for spec in "$@"; do
  name=${spec%%:*}; flags=${spec#*:}
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --fmad=true \
      -Xcompiler -fPIC,-O2,-ffp-contract=off -shared -o ../../build/lib_$name.so *.cu $flags &
done
wait
ls -la ../../build/
