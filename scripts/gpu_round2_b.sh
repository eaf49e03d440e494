#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/b_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/b_pytest.log
tail -15 gpurun_out/b_pytest.log
for v in 6 7 0 1; do
  timeout 200 python bench.py --N 512x512x64 --bc ddp --pair-variant $v --no-e2e > gpurun_out/b_proxy_v$v.json 2> gpurun_out/b_proxy_v$v.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/b_proxy_v$v.json"))
    print("proxy v$v", d["ms_per_step"], d["value"], d.get("validated_vs_oracle"), d["plan"][-120:])
except Exception as e:
    print("proxy v$v failed", e)
PY
done
# ncu of the pair kernel on the proxy block (one-wave, all CTAs are face CTAs)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pair_step -s 6 -c 1 -o gpurun_out/b_ncu_proxy_v4 -f \
  python bench.py --N 512x512x64 --bc ddp --pair-variant 4 --no-e2e --no-graph --reps 1 > gpurun_out/b_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
