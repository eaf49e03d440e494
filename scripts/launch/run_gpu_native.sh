#!/bin/bash
# Python-free variant: one process, G GPUs, peer access (csrc/tools/mpigpu_main.cpp)
# usage: run_gpu_native.sh [G | PxQxR] [N tau K ...]     G: z slabs over G GPUs; PxQxR: a grid of 3-D blocks
cd "$(dirname "$0")/../.." && make -s cuda tools
G=${1:-1}; shift || true
case "$G" in
  *x*) IFS=x read -r P Q R <<< "$G"; bin/mpigpu --dims "$P" "$Q" "$R" ${*:-512 0.001 20 1} ;;
  *)   bin/mpigpu -np "$G" ${*:-512 0.001 20 1} ;;
esac
