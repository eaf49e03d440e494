#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "streaming or pair_mode or exchange_stat or checkpoint" > gpurun_out/j_pytest.log 2>&1; tail -2 gpurun_out/j_pytest.log
# per-GPU blocks of 1024^3 on 8 GPUs: slab 1x1x8 (wide halo) vs block 1x2x4 (fix-up protocol)
bash scripts/gpu_variants.sh j "default" "1024x1024x128" --steps 100 --L pi --no-nccl-baseline
bash scripts/gpu_variants.sh jd "default" "1024x1024x128" --steps 100 --L pi --bc ddd --no-nccl-baseline
timeout 200 python bench.py --N 1024x512x256 --bc dpp --L pi --steps 100 --no-e2e --no-nccl-baseline > gpurun_out/j_blk.json 2>gpurun_out/j_blk.err; python -c "
import json; d=json.load(open('gpurun_out/j_blk.json')); print('block proxy 1023x512x256 dpp', d['ms_per_step'], d['value'], d['validated_vs_oracle'], d['exchange'], d['plan'][-150:])"
timeout 200 python bench.py --N 2048x2048x256 --bc ddp --L 4 --steps 20 --reps 3 --no-e2e --no-nccl-baseline > gpurun_out/j_2048slab.json 2>gpurun_out/j_2048slab.err; python -c "
import json; d=json.load(open('gpurun_out/j_2048slab.json')); print('slab proxy 2047x2047x256', d['ms_per_step'], d['value'], d['validated_vs_oracle'], d['exchange'], d['plan'][-150:])"
timeout 200 python bench.py --N 2048x1024x512 --bc dpp --L 4 --steps 20 --reps 3 --no-e2e --no-nccl-baseline > gpurun_out/j_2048blk.json 2>gpurun_out/j_2048blk.err; python -c "
import json; d=json.load(open('gpurun_out/j_2048blk.json')); print('block proxy 2047x1024x512', d['ms_per_step'], d['value'], d['validated_vs_oracle'], d['exchange'], d['plan'][-150:])"
python bench.py --no-nccl-baseline > gpurun_out/j_512.json 2> gpurun_out/j_512.err; python -c "
import json; d=json.load(open('gpurun_out/j_512.json')); print('512 1gpu', d['ms_per_step'], d['value'], 'e2e', d['e2e']['value'], d['e2e']['wall_s'], 'total_s', d['total_s'], d['init_s'])"
