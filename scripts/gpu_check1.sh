#!/bin/bash
# one GPU: the whole GPU test suite, the smoke entry and the default bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q ${PYTEST_ARGS} > gpurun_out/check_pytest_1gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/check_pytest_1gpu.log
tail -${TAILN:-5} gpurun_out/check_pytest_1gpu.log
[ -n "$SKIP_BENCH" ] && exit 0
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 300 python bench.py > gpurun_out/check_bench_512.json 2> gpurun_out/check_bench_512.err
python -c "
import json; d=json.load(open('gpurun_out/check_bench_512.json')); print('bench', d['ms_per_step'], d['value'], d['validated_vs_oracle'], 'e2e', d['e2e']['value'], 'nccl', d['baseline_nccl'].get('ms_per_step'), d['vs_nccl'], 'total_s', d['total_s'], 'vs_baseline', d['vs_baseline'], 'launches', d['gpu_launches'], d['clocks'])"
python bench.py --impl reference
