#!/bin/bash
# report.pdf p.15: #BSUB -gpu "num=G:mode=exclusive_process"; mpirun -np G ./mpigpu-1 512 0.001 20 1
# usage: run_gpu.sh [G] [N tau K ...]   (one process per GPU over torch.distributed/NCCL)
cd "$(dirname "$0")/../.."
G=${1:-1}; shift || true
ARGS=${*:-512 0.001 20 1}
if [ "$G" = "1" ]; then
  python -m mpi_cuda_b200 $ARGS --backend fused
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node "$G" --master-addr 127.0.0.1 --master-port ${PORT:-29555} \
    -m mpi_cuda_b200 $ARGS --backend fused --decomp auto
fi
