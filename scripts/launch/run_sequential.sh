#!/bin/bash
# report.pdf p.12, p.20: export OMP_NUM_THREADS=1; for grid in 128 256 512: ./wave $grid 0.001 20
cd "$(dirname "$0")/../.." && make -s cpu
export OMP_NUM_THREADS=1
for grid in ${GRIDS:-128 256 512}; do
  bin/wave "$grid" 0.001 20
done
