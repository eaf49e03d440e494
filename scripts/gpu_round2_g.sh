#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "pair or wide or graph or golden or streaming" > gpurun_out/g_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/g_pytest.log
tail -3 gpurun_out/g_pytest.log
echo "--- PDL on"
bash scripts/gpu_variants.sh g_pdl1 "default" "512x512x64 513x513x64 513x512x64 512x513x64 512x512x128 512" --pair-variant 4
echo "--- PDL off"
W3D_PDL=0 bash scripts/gpu_variants.sh g_pdl0 "default" "512x512x64 513x513x64 512" --pair-variant 4
