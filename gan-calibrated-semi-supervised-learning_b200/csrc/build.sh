#!/bin/bash
# Builds libb2c.so (the C-ABI library, include/b2c.h) for sm_100a, in-tree next to the Python package.
set -e
cd "$(dirname "$0")"
OUT=../b200cgan/libb2c.so
SRCS="host.cu tapgemm.cu convt64.cu wgrad.cu pack.cu first.cu head.cu norm.cu sn.cu loss.cu optim.cu refine.cu simple.cu dataset.cu"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=default"
mkdir -p build
pids=()
for s in $SRCS; do
  o=build/${s%.cu}.o
  if [ ! -f $o ] || [ $s -nt $o ] || [ common.cuh -nt $o ] || [ tmap.h -nt $o ]; then
    $NVCC $FLAGS -c $s -o $o &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -o $OUT $(for s in $SRCS; do echo build/${s%.cu}.o; done) -lcudart
echo "built $OUT"
