#!/bin/bash
mkdir -p gpurun_out
for n in 512x512x64 513x513x64 512x512x128 513x513x128; do
  timeout 200 python bench.py --N $n --bc ddp --pair-variant 4 --no-e2e > gpurun_out/c_proxy_$n.json 2> gpurun_out/c_proxy_$n.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/c_proxy_$n.json"))
    print("proxy $n", round(d["ms_per_step"],5), round(d["value"],1), d.get("validated_vs_oracle"), d["exchange"])
except Exception as e:
    print("proxy $n failed", e)
PY
done
