#!/bin/bash
# report.pdf p.13-14, p.25-26: export OMP_NUM_THREADS=t; mpiexec ./mpiomp $grid 0.001 20 t
cd "$(dirname "$0")/../.." && make -s cpu
T=${OMP_NUM_THREADS:-4}
for grid in ${GRIDS:-128 256 512}; do
  for P in ${RANKS:-1 2}; do
    OMP_NUM_THREADS=$T bin/mpiomp -np "$P" "$grid" 0.001 20 "$T" --quiet
  done
done
