#!/bin/bash
mkdir -p gpurun_out
python scripts/timeline.py --N 512x512x64 --bc ddp --pair-variant 4 > gpurun_out/h_timeline_graph.txt 2>&1; tail -14 gpurun_out/h_timeline_graph.txt
W3D_PDL=1 python scripts/timeline.py --N 512x512x64 --bc ddp --pair-variant 4 > gpurun_out/h_timeline_pdl.txt 2>&1; tail -5 gpurun_out/h_timeline_pdl.txt
python scripts/timeline.py --N 512x512x64 --bc ddp --pair-variant 4 --no-graph > gpurun_out/h_timeline_eager.txt 2>&1; tail -3 gpurun_out/h_timeline_eager.txt
python scripts/timeline.py --N 512 > gpurun_out/h_timeline_512.txt 2>&1; tail -3 gpurun_out/h_timeline_512.txt
