#!/bin/bash
TAG=${1:-e8}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -q -k "native" > gpurun_out/${TAG}_pytest_native.log 2>&1; tail -2 gpurun_out/${TAG}_pytest_native.log
for N in 8 4 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 3 --no-nccl-baseline > gpurun_out/${TAG}_g$N.json 2> gpurun_out/${TAG}_g$N.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/${TAG}_g$N.json") if l.startswith("{")][-1])
    print("g$N: %.5f ms/step %.1f GLUP/s valid=%s e2e=%.1f (%.3f ms) loop %.3f ms total %.3f ms" % (d["ms_per_step"], d["value"], d.get("validated_vs_oracle"), d["e2e"]["value"], d["e2e"]["wall_s"]*1e3, d["loop_s"]*1e3, d["total_s"]*1e3))
except Exception as e:
    print("g$N failed", e, open("gpurun_out/${TAG}_g$N.err").read()[-800:])
PY
done
