#!/bin/bash
# usage: gpu_multi.sh TAG NGPU [run tests: 0/1]
TAG=$1; N=$2; TESTS=${3:-1}
mkdir -p gpurun_out
nvidia-smi -L | wc -l
if [ "$TESTS" = "1" ]; then
  timeout 1500 python -m pytest tests/test_gpu_multi.py -m gpu -q > gpurun_out/${TAG}_pytest_multi.log 2>&1
  echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_multi.log
  tail -25 gpurun_out/${TAG}_pytest_multi.log
fi
run() { # name, extra args
  name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/${TAG}_$name.json") if l.startswith("{")][-1])
    x = d.get("exchange", {})
    print("$name: %.5f ms/step %.1f GLUP/s valid=%s e2e=%s wait %.2f copy %.2f us/pass | %s" % (d["ms_per_step"], d["value"], d.get("validated_vs_oracle"), (d.get("e2e") or {}).get("value"), 2e3*x.get("flag_wait_ms_per_step",0), 2e3*x.get("halo_copy_ms_per_step",0), d.get("plan","").split("||")[-1][:90]))
except Exception as e:
    print("$name failed", e, open("gpurun_out/${TAG}_$name.err").read()[-800:])
PY
}
run 512 --steps 20 --warmup 3
run 512_nopair --steps 20 --warmup 3 --no-pair --no-e2e
run 1024 --N 1024 --L pi --steps 100 --no-e2e
