#!/bin/bash
# round-2 first GPU pass: tests, headline bench, proxy sweep of the 8-GPU slab block on one GPU
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/a_gpus.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
timeout 300 python bench.py > gpurun_out/a_bench_512.json 2> gpurun_out/a_bench_512.err
tail -c 600 gpurun_out/a_bench_512.json
for v in 4 8 9 10 3; do
  timeout 200 python bench.py --N 512x512x64 --bc ddp --pair-variant $v --no-e2e > gpurun_out/a_proxy_v$v.json 2> gpurun_out/a_proxy_v$v.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/a_proxy_v$v.json"))
    print("proxy v$v", d["ms_per_step"], d["value"], d.get("validated_vs_oracle"), d["plan"][-120:])
except Exception as e:
    print("proxy v$v failed", e)
PY
done
