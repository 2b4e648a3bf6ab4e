#!/bin/bash
# round 2, pass m: does warming L2 at the start of a frame pay when the frame STARTS COLD (bench.py flushes L2 before
# every step)?  whole C4 frame and rank 0's share of an 8-rank frame; prefetches (1) and real loads (2)
tag=${1:-r2m}
mkdir -p gpurun_out
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | cut -c1-70)" >> $out; }
for rep in 1 2; do
  for warm in 0 1 2; do
    t "cold L2, warm=$warm c4 whole" CRT_FLUSH=1 CRT_WARM_L2=$warm timeout 200 python tools/render_frames.py c4 8 auto kd 0
    t "cold L2, warm=$warm c4 rank0of8" CRT_FLUSH=1 CRT_WARM_L2=$warm CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 8 auto kd 0
  done
  t "warm L2 (no flush) c4 rank0of8" CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 8 auto kd 0
done
cat $out
