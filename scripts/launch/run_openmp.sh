#!/bin/bash
# report.pdf p.21 §5.2.4: for threads in 1 2 4 8 16 32 64: ./openmpwave $size 0.001 20 $threads
cd "$(dirname "$0")/../.." && make -s cpu
for size in ${GRIDS:-128 256 512}; do
  for threads in ${THREADS:-1 2 4 8}; do
    echo "== $size^3, $threads threads"
    bin/openmpwave "$size" 0.001 20 "$threads" --quiet
  done
done
