#!/bin/bash
# builds experimental variants of the CUDA library next to the default one: libwave3d_b200.<tag>.so
# usage: build_variants.sh tag "-DFOO=1 -DBAR=0" [tag2 "..."]...
set -e
cd "$(dirname "$0")/.."
while [ $# -ge 2 ]; do
  tag=$1; defs=$2; shift 2
  nvcc -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Iinclude \
       -Xcompiler -fPIC,-fopenmp $defs -shared -o mpi_cuda_b200/_native/libwave3d_b200.$tag.so csrc/cuda/*.cu -cudart static &
done
wait
ls -la mpi_cuda_b200/_native/
