#!/bin/bash
# Run the multi-GPU measurement suite on the box gpurun gives us.  usage: scripts/gpu_suite.sh NGPUS [tag]
# Writes one JSON line per run into gpurun_out/suite_<tag>_<name>.json (and .log on failure).
N=${1:-2}; TAG=${2:-g$N}; PORT=29600
mkdir -p gpurun_out
run() {  # name, extra bench args...
  local name=$1; shift
  PORT=$((PORT+1))
  if [ "$N" = "1" ]; then
    timeout 300 python bench.py --gpus 1 "$@" > gpurun_out/suite_${TAG}_${name}.json 2> gpurun_out/suite_${TAG}_${name}.log
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $N "$@" > gpurun_out/suite_${TAG}_${name}.json 2> gpurun_out/suite_${TAG}_${name}.log
  fi
  echo "$name rc=$? $(tail -c 300 gpurun_out/suite_${TAG}_${name}.json | tr -d '\n' | head -c 10)"
}
# SUITE_SET=pair: only the temporal-blocking runs (names *_pair_*); default: single-pass fused + NCCL baseline
if [ "${SUITE_SET:-}" = "pair" ]; then
  run 512_pair_block --steps 20 --warmup 3
  [ "$N" != "1" ] && run 512_pair_slab --steps 20 --warmup 3 --decomp slab --no-e2e
  run 1024_pair_block --N 1024 --L pi --steps 100 --warmup 3 --reps 3 --no-e2e
  case $N in
    1) WN=512x512x512;   WL=0.5pix0.5pix0.5pi;;
    2) WN=512x512x1024;  WL=0.5pix0.5pixpi;;
    4) WN=512x1024x1024; WL=0.5pixpixpi;;
    8) WN=1024x1024x1024; WL=pixpixpi;;
  esac
  [ "$N" != "8" ] && run weak_pair_block --N $WN --L $WL --steps 100 --warmup 3 --reps 3 --no-e2e
  [ "$N" = "8" ] && run 2048_pair_block --N 2048 --L 4 --bc ddp --steps 1000 --warmup 3 --reps 1 --no-e2e
  exit 0
fi
# headline: 512^3 K=20 (strong scaling)
run 512_fused_block --steps 20 --warmup 3 --no-pair
run 512_fused_slab  --steps 20 --warmup 3 --decomp slab --no-e2e --no-pair
run 512_nccl_block  --steps 20 --warmup 3 --backend nccl
# 1024^3 K=100, L=pi (CFL-stable), strong scaling, 3-D blocks and slabs
run 1024_fused_block --N 1024 --L pi --steps 100 --warmup 3 --reps 3 --no-e2e --no-pair
run 1024_fused_slab  --N 1024 --L pi --steps 100 --warmup 3 --reps 3 --no-e2e --decomp slab --no-pair
run 1024_nccl_block  --N 1024 --L pi --steps 100 --warmup 3 --reps 3 --backend nccl
# weak scaling: 512^3 unknowns per GPU (grid grows along z, then y)
case $N in
  1) WN=512x512x512;   WL=0.5pix0.5pix0.5pi;;
  2) WN=512x512x1024;  WL=0.5pix0.5pixpi;;
  4) WN=512x1024x1024; WL=0.5pixpixpi;;
  8) WN=1024x1024x1024; WL=pixpixpi;;
esac
run weak_fused_block --N $WN --L $WL --steps 100 --warmup 3 --reps 3 --no-e2e --no-pair
run weak_nccl_block  --N $WN --L $WL --steps 100 --warmup 3 --reps 3 --backend nccl
