#!/bin/bash
# final evidence on one GPU: full GPU test suite, headline bench, launch list, ncu --set full of every hot kernel
TAG=${1:-p}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_1gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_1gpu.log
tail -4 gpurun_out/${TAG}_pytest_1gpu.log
timeout 300 python bench.py > gpurun_out/${TAG}_bench_512.json 2> gpurun_out/${TAG}_bench_512.err
python -c "
import json; d=json.load(open('gpurun_out/${TAG}_bench_512.json')); print('bench', d['ms_per_step'], d['value'], d['validated_vs_oracle'], 'e2e', d['e2e']['value'], 'nccl', d['baseline_nccl'].get('ms_per_step'), d['vs_nccl'], 'total_s', d['total_s'], 'vs_baseline', d['vs_baseline'])"
NC="--set full --clock-control none --import-source on"
B="python bench.py --no-e2e --no-nccl-baseline --no-graph --reps 1"
# launch list of the default path (no metrics but the duration)
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_launches_512.csv $B > /dev/null 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_launches_slab8.csv $B --N 512x512x64 --bc ddp > /dev/null 2>&1
timeout 400 ncu $NC -k regex:pair_step -s 4 -c 1 -o gpurun_out/${TAG}_ncu_pair_512 -f $B > gpurun_out/${TAG}_ncu1.log 2>&1
timeout 400 ncu $NC -k regex:pair_step -s 4 -c 1 -o gpurun_out/${TAG}_ncu_pair_slab8 -f $B --N 512x512x64 --bc ddp > gpurun_out/${TAG}_ncu2.log 2>&1
timeout 400 ncu $NC -k regex:pair_step -s 4 -c 1 -o gpurun_out/${TAG}_ncu_pair_cluster -f $B --N 512x512x64 --bc ddp --pair-variant 11 > gpurun_out/${TAG}_ncu3.log 2>&1
timeout 400 ncu $NC -k regex:fixup -s 4 -c 1 -o gpurun_out/${TAG}_ncu_fixup -f $B --N 512x512x64 --bc dpp > gpurun_out/${TAG}_ncu4.log 2>&1
timeout 400 ncu $NC -k regex:fused_step -s 4 -c 1 -o gpurun_out/${TAG}_ncu_fused_512 -f $B --no-pair > gpurun_out/${TAG}_ncu5.log 2>&1
ls -la gpurun_out/${TAG}_*.ncu-rep
