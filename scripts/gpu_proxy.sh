#!/bin/bash
# usage: gpu_proxy.sh TAG [pytest-filter]  — one-GPU stand-ins for the per-GPU blocks of multi-GPU slab runs
# (z periodic: the block exchanges halos with itself through the same in-kernel protocol)
TAG=$1; FILTER=$2
mkdir -p gpurun_out
if [ -n "$FILTER" ]; then
  timeout 1500 python -m pytest tests -m gpu -q -x -k "$FILTER" > gpurun_out/${TAG}_pytest.log 2>&1
  echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
  tail -4 gpurun_out/${TAG}_pytest.log
fi
for n in 512x512x64 513x513x64 512x512x128 512x512x256 512; do
  bc=ddp; [ "$n" = "512" ] && bc=ddd
  timeout 200 python bench.py --N $n --bc $bc --no-e2e ${PROXY_ARGS} > gpurun_out/${TAG}_proxy_$n.json 2> gpurun_out/${TAG}_proxy_$n.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_proxy_$n.json"))
    x = d["exchange"]
    print("proxy $n %.5f ms/step %.1f GLUP/s valid=%s wait %.2f us/pass copy %.2f us/pass | %s" % (d["ms_per_step"], d["value"], d.get("validated_vs_oracle"), 2e3*x["flag_wait_ms_per_step"], 2e3*x["halo_copy_ms_per_step"], d["plan"].split("||")[1][:70]))
except Exception as e:
    print("proxy $n failed", e, open("gpurun_out/${TAG}_proxy_$n.err").read()[-600:])
PY
done
