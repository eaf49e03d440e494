#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/i_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/i_pytest.log
tail -3 gpurun_out/i_pytest.log
bash scripts/gpu_variants.sh i "default" "512x512x64 513x513x64 512x512x128 512x512x256 512" --pair-variant 4
