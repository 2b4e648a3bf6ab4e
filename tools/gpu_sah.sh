#!/bin/bash
# SAH cost sweep of the walk-tree builder with the lean kernels (latency-bound: do fewer, fatter leaves pay now?)
tag=${1:-sah}
mkdir -p gpurun_out
out=gpurun_out/${tag}_sah.txt
: > $out
for cfg in "1 1 1.5 1.0" "1 1 1.0 1.0" "1 1 0.6 1.0" "1 0.5 0.6 1.0" "1 0.5 0.3 1.0" "1 0.3 0.15 1.0" "1 0.2 0.08 1.0" "1 2 2.8 1.0" "1.5 1 1.5 1.0" "2 1 1.5 1.0" "3 1 1.5 1.0"; do
  set -- $cfg
  export CRT_SAH_NODE=$1 CRT_SAH_LEAF=$2 CRT_SAH_TRI=$3 CRT_SAH_EMPTY=$4
  a=$(timeout 120 python tools/render_frames.py c4 5 auto kd 0 | tail -1 | grep -o "'gpu_ms': [0-9.]*")
  b=$(CRT_RANK=0 CRT_WORLD=8 timeout 120 python tools/render_frames.py c4 5 auto kd 0 | tail -1 | grep -o "'gpu_ms': [0-9.]*")
  c=$(timeout 120 python tools/render_frames.py c4 1 auto kd 1 | tail -1 | grep -o "'node_tests.*")
  g=$(timeout 200 python tools/render_frames.py c3-gi 2 auto kd 0 | tail -1 | grep -o "'gpu_ms': [0-9.]*")
  echo "node=$1 leaf=$2 tri=$3: whole $a | rank0of8 $b | gi $g | $c" | tee -a $out
done
