#!/bin/bash
# 8-GPU box: all multi-GPU tests (2, 4, 8 ranks) and the 3-D block decompositions with the edge-line wide halo
TAG=${1:-b8}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -v > gpurun_out/${TAG}_pytest_multi.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_multi.log
tail -4 gpurun_out/${TAG}_pytest_multi.log
grep -c PASSED gpurun_out/${TAG}_pytest_multi.log
run() { name=$1; shift
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --no-e2e --no-nccl-baseline "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/${TAG}_$name.json") if l.startswith("{")][-1])
    x = d.get("exchange", {})
    print("$name: %.5f ms/step %.1f GLUP/s valid=%s wait %.2f copy %.2f us/pass | %s" % (d["ms_per_step"], d["value"], d.get("validated_vs_oracle"), 2e3*x.get("flag_wait_ms_per_step",0), 2e3*x.get("halo_copy_ms_per_step",0), d.get("plan","").split("||")[-1][-110:]))
except Exception as e:
    print("$name failed", e, open("gpurun_out/${TAG}_$name.err").read()[-800:])
PY
}
run 1024_1x2x4_wide --N 1024 --L pi --steps 100 --decomp 1x2x4
run 1024_2x2x2_wide --N 1024 --L pi --steps 100 --decomp 2x2x2
run 1024_slab --N 1024 --L pi --steps 100
run 512_1x2x4_wide --steps 20 --decomp 1x2x4
run 512_1x2x4_fixup --steps 20 --decomp 1x2x4 --fixup
run 512_slab --steps 20
