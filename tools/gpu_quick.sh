#!/bin/bash
# quick timing pass of the current build on every workload.  usage: tools/gpu_quick.sh <tag>
tag=${1:-quick}
mkdir -p gpurun_out
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | cut -c1-100)" >> $out; }
for rep in 1 2; do
  for w in c4 c1 c3 c5; do t "$w" timeout 200 python tools/render_frames.py $w 6 auto kd 0; done
  t "c4 rank0of8" CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 auto kd 0
done
t "c3-gi" timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
cat $out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${tag}_launches_c4.csv python tools/render_frames.py c4 3 auto kd 0 > /dev/null 2>&1
