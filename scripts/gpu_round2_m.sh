#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "streaming or native or checkpoint or small_and_odd" > gpurun_out/m_pytest.log 2>&1; tail -4 gpurun_out/m_pytest.log
python bench.py --no-nccl-baseline > gpurun_out/m_512.json 2> gpurun_out/m_512.err; python -c "
import json; d=json.load(open('gpurun_out/m_512.json')); print('512 1gpu', d['ms_per_step'], d['value'], 'e2e', d['e2e']['value'], d['e2e']['wall_s'], 'total_s', d['total_s'])" || tail -5 gpurun_out/m_512.err
python bench.py --no-nccl-baseline --N 512x512x64 --bc ddp > gpurun_out/m_slab.json 2> gpurun_out/m_slab.err; python -c "
import json; d=json.load(open('gpurun_out/m_slab.json')); print('slab8 proxy', d['ms_per_step'], d['value'], 'e2e', d['e2e']['value'], d['e2e']['wall_s'], 'loop_s', d['loop_s'])" || tail -5 gpurun_out/m_slab.err
