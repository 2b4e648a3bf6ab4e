#!/bin/bash
# per-rank frame time of an 8-rank C4 frame for every rank and a few band heights (the barrier waits for the slowest rank)
tag=${1:-band}
mkdir -p gpurun_out
out=gpurun_out/${tag}_bands.txt
: > $out
for band in 8 4 16 32; do
  line="band_rows=$band:"
  for r in 0 1 2 3 4 5 6 7; do
    ms=$(CRT_BAND=$band CRT_RANK=$r CRT_WORLD=8 timeout 120 python tools/render_frames.py c4 5 auto kd 0 | tail -1 | grep -o "'gpu_ms': [0-9.]*" | cut -c11-16)
    line="$line $ms"
  done
  echo "$line" | tee -a $out
done
