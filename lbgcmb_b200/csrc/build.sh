#!/bin/bash
# Build liblbgcmb.so for sm_100a (in-tree, next to the python package).  Usage: build.sh [extra nvcc flags]
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
OUT=../liblbgcmb.so
SRCS=$(ls *.cu)
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --fmad=true \
      -Xcompiler -fPIC,-O2,-ffp-contract=off -shared -o $OUT $SRCS "$@"
echo "built $(readlink -f $OUT)"
