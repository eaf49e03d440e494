#!/bin/bash
# Python-free variant: one process, G GPUs, peer access (csrc/tools/mpigpu_main.cpp)
cd "$(dirname "$0")/../.." && make -s cuda tools
G=${1:-1}; shift || true
bin/mpigpu -np "$G" ${*:-512 0.001 20 1}
