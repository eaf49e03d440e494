for a in "" "--fixup"; do
timeout 200 python bench.py --N 1024x512x256 --bc dpp --L pi --steps 100 --no-e2e --no-nccl-baseline $a > gpurun_out/q_blk$a.json 2>gpurun_out/q_blk.err; python -c "
import json; d=json.load(open('gpurun_out/q_blk$a.json')); print('block proxy 1023x512x256 dpp $a', round(d['ms_per_step'],4), round(d['value'],1), d['validated_vs_oracle'], {k:round(v,3) for k,v in d['exchange'].items() if 'per_step' in k}, d['plan'][-90:])"
done
timeout 200 python bench.py --N 1024x1024x128 --bc ddp --L pi --steps 100 --no-e2e --no-nccl-baseline > gpurun_out/q_slab.json 2>gpurun_out/q_slab.err; python -c "
import json; d=json.load(open('gpurun_out/q_slab.json')); print('slab proxy 1023x1023x128 ddp', round(d['ms_per_step'],4), round(d['value'],1), d['validated_vs_oracle'])"
for a in "" "--fixup"; do
timeout 200 python bench.py --N 512x256x128 --bc dpp --steps 20 --no-e2e --no-nccl-baseline $a > gpurun_out/q_blk512$a.json 2>gpurun_out/q_blk.err; python -c "
import json; d=json.load(open('gpurun_out/q_blk512$a.json')); print('block proxy 511x256x128 dpp $a', round(d['ms_per_step'],4), round(d['value'],1), d['validated_vs_oracle'], d['plan'][-90:])"
done
