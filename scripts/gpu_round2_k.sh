#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "cluster" > gpurun_out/k_pytest.log 2>&1; tail -4 gpurun_out/k_pytest.log
bash scripts/gpu_variants.sh k11 "default" "512x512x64 513x513x64 512" --pair-variant 11 --no-nccl-baseline
bash scripts/gpu_variants.sh k4 "default" "512x512x64 513x513x64 512" --pair-variant 4 --no-nccl-baseline
