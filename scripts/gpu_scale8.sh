#!/bin/bash
# one 8-GPU box: multi-GPU tests, the 1/2/4/8 scaling table at 512^3, weak scaling, the large runs,
# per-rank device timelines, NVLink byte counters around a long native run, the CLI under torchrun
TAG=${1:-s8}
mkdir -p gpurun_out
nvidia-smi -L | wc -l
nvidia-smi topo -m > gpurun_out/${TAG}_topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -v > gpurun_out/${TAG}_pytest_multi.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_multi.log
tail -4 gpurun_out/${TAG}_pytest_multi.log
run() { # name, ngpu, extra args
  name=$1; N=$2; shift 2
  if [ "$N" = "1" ]; then
    timeout 300 python bench.py --gpus 1 "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  fi
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/${TAG}_$name.json") if l.startswith("{")][-1])
    x = d.get("exchange", {}); b = d.get("baseline_nccl") or {}
    print("$name: %.5f ms/step %.1f GLUP/s valid=%s e2e=%s nccl=%s wait %.2f copy %.2f us/pass roofline %.2f | %s" % (d["ms_per_step"], d["value"], d.get("validated_vs_oracle"), (d.get("e2e") or {}).get("value"), b.get("ms_per_step", b.get("error")), 2e3*x.get("flag_wait_ms_per_step",0), 2e3*x.get("halo_copy_ms_per_step",0), d["roofline"]["fraction_of_roofline"], d.get("plan","").split("||")[-1][:100]))
except Exception as e:
    print("$name failed", e, open("gpurun_out/${TAG}_$name.err").read()[-800:])
PY
}
run g8_512 8 --steps 20 --warmup 3
run g4_512 4 --steps 20 --warmup 3
run g2_512 2 --steps 20 --warmup 3
run g1_512 1 --steps 20 --warmup 3
run g8_1024 8 --N 1024 --L pi --steps 100 --no-e2e
run g8_1024_block 8 --N 1024 --L pi --steps 100 --no-e2e --no-nccl-baseline --decomp 1x2x4
run g1_weak 1 --N 1024x1024x128 --L pi --steps 100 --no-e2e --no-nccl-baseline
run g8_2048 8 --N 2048 --L 4 --bc ddp --steps 1000 --reps 3 --no-e2e --no-nccl-baseline
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29534 scripts/timeline.py --N 512 --steps 20 > gpurun_out/${TAG}_timeline8.txt 2>&1
grep "launches" gpurun_out/${TAG}_timeline8.txt | head -2
python scripts/nvlink_bytes.py > gpurun_out/${TAG}_nvl0.json 2> gpurun_out/${TAG}_nvl.err
timeout 200 bin/mpigpu -np 8 512 0.001 2000 --quiet > gpurun_out/${TAG}_mpigpu8_2000.txt 2>&1; tail -2 gpurun_out/${TAG}_mpigpu8_2000.txt
python scripts/nvlink_bytes.py > gpurun_out/${TAG}_nvl1.json 2>> gpurun_out/${TAG}_nvl.err
python scripts/nvlink_bytes.py gpurun_out/${TAG}_nvl0.json gpurun_out/${TAG}_nvl1.json > gpurun_out/${TAG}_nvlink_delta.txt 2>&1; head -3 gpurun_out/${TAG}_nvlink_delta.txt
timeout 200 bin/mpigpu --dims 1 2 4 512 0.001 20 > gpurun_out/${TAG}_mpigpu_1x2x4.txt 2>&1; tail -2 gpurun_out/${TAG}_mpigpu_1x2x4.txt
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29535 -m mpi_cuda_b200 512 0.001 20 > gpurun_out/${TAG}_cli8.txt 2>&1; tail -5 gpurun_out/${TAG}_cli8.txt
