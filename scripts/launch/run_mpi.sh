#!/bin/bash
# report.pdf p.12-13, p.22-24: mpirun -np P ./mpi $grid 0.001 20, P in {1,2,4,8,10,16,20,32}
# Ranks are threads of one process here (no MPI on the image); for one OS process per rank use
#   torchrun --nnodes=1 --nproc-per-node P --master-addr 127.0.0.1 -m mpi_cuda_b200 $grid 0.001 20 --backend gloo
cd "$(dirname "$0")/../.." && make -s cpu
for grid in ${GRIDS:-128 256 512}; do
  for P in ${RANKS:-1 2 4 8}; do
    bin/mpi -np "$P" "$grid" 0.001 20 --quiet
  done
done
