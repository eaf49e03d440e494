#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -k "native" > gpurun_out/l_pytest.log 2>&1; tail -5 gpurun_out/l_pytest.log
timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -m gpu -q -k "native or noalloc" > gpurun_out/l_pytest1.log 2>&1; tail -5 gpurun_out/l_pytest1.log
nvidia-smi nvlink -gt d -i 0 2>&1 | head -8
timeout 120 bin/mpigpu -np 2 512 0.001 200 --quiet
nvidia-smi nvlink -gt d -i 0 2>&1 | head -8
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 -m mpi_cuda_b200 512 0.001 20 2>&1 | tail -8
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 -m mpi_cuda_b200 256 0.001 20 --trace gpurun_out/l_trace.json 2>&1 | tail -3
