#!/bin/bash
# report.pdf p.12-13, p.22-24: mpirun -np P ./mpi $grid 0.001 20, P in {1,2,4,8,10,16,20,32}
#
# Two equivalents here (the image has no MPI):
#   1. bin/mpi -np P ...: the native program, ranks are threads of ONE process that exchange halos
#      through an in-memory mailbox (csrc/cpu/wave_cpu.cpp);
#   2. MODE=procs: ONE OS PROCESS PER RANK, the closest thing to `mpirun -np P`: torchrun starts P
#      Python processes that exchange halos over torch.distributed / gloo (TCP on 127.0.0.1) and
#      reduce the errors with all_reduce, each rank stepping its block with the same native kernels.
# MODE=threads (default) runs 1, MODE=procs runs 2, MODE=both runs both and the traces can be compared.
cd "$(dirname "$0")/../.." && make -s cpu
MODE=${MODE:-threads}
for grid in ${GRIDS:-128 256 512}; do
  for P in ${RANKS:-1 2 4 8}; do
    if [ "$MODE" = "threads" ] || [ "$MODE" = "both" ]; then
      bin/mpi -np "$P" "$grid" 0.001 20 --quiet
    fi
    if [ "$MODE" = "procs" ] || [ "$MODE" = "both" ]; then
      if [ "$P" = "1" ]; then
        python -m mpi_cuda_b200 "$grid" 0.001 20 --backend gloo
      else
        python -m torch.distributed.run --nnodes=1 --nproc-per-node "$P" --master-addr 127.0.0.1 \
          --master-port "${MASTER_PORT:-29517}" -m mpi_cuda_b200 "$grid" 0.001 20 --backend gloo
      fi
    fi
  done
done
