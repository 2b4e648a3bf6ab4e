#!/bin/bash
tag=${1:-ipl}
mkdir -p gpurun_out
out=gpurun_out/${tag}_items_per_lane.txt
: > $out
for k in 1 2 3 4 6 8; do
  a=$(CRT_ITEMS_PER_LANE=$k timeout 120 python tools/render_frames.py c4 6 auto kd 0 | tail -1 | grep -o "'gpu_ms': [0-9.]*" | cut -c11-16)
  b=$(CRT_ITEMS_PER_LANE=$k CRT_RANK=0 CRT_WORLD=8 timeout 120 python tools/render_frames.py c4 6 auto kd 0 | tail -1 | grep -o "'gpu_ms': [0-9.]*" | cut -c11-16)
  b3=$(CRT_ITEMS_PER_LANE=$k CRT_RANK=3 CRT_WORLD=8 timeout 120 python tools/render_frames.py c4 6 auto kd 0 | tail -1 | grep -o "'gpu_ms': [0-9.]*" | cut -c11-16)
  c=$(CRT_ITEMS_PER_LANE=$k CRT_RANK=0 CRT_WORLD=2 timeout 120 python tools/render_frames.py c4 6 auto kd 0 | tail -1 | grep -o "'gpu_ms': [0-9.]*" | cut -c11-16)
  d=$(CRT_ITEMS_PER_LANE=$k timeout 120 python tools/render_frames.py c3 6 auto kd 0 | tail -1 | grep -o "'gpu_ms': [0-9.]*" | cut -c11-16)
  e=$(CRT_ITEMS_PER_LANE=$k timeout 120 python tools/render_frames.py c5 6 auto kd 0 | tail -1 | grep -o "'gpu_ms': [0-9.]*" | cut -c11-16)
  echo "items_per_lane=$k: c4 whole $a | rank0of8 $b rank3of8 $b3 | rank0of2 $c | c3 $d | c5 $e" | tee -a $out
done
