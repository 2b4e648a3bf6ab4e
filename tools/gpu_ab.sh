#!/bin/bash
# A/B of the default build against the variants given: tools/gpu_ab.sh <tag> <variants...>   (two passes, same box)
tag=${1:-ab}; shift
variants=$@
mkdir -p gpurun_out
V=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | grep -o "'gpu_ms': [0-9.]*")" >> $out; }
for rep in 1 2; do
  for v in default $variants; do
    lib=$V/libcrt_b200_$v.so; [ $v = default ] && lib=$PWD/chaos-ray-tracing-course-2025_b200/lib/libcrt_b200.so
    t "$v c4" CRT_B200_LIB=$lib timeout 200 python tools/render_frames.py c4 6 auto kd 0
    t "$v c4 rank0of8" CRT_B200_LIB=$lib CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 auto kd 0
    t "$v c1" CRT_B200_LIB=$lib timeout 200 python tools/render_frames.py c1 8 auto kd 0
    t "$v c3" CRT_B200_LIB=$lib timeout 200 python tools/render_frames.py c3 8 auto kd 0
    [ $rep = 1 ] && t "$v c5" CRT_B200_LIB=$lib timeout 200 python tools/render_frames.py c5 6 auto kd 0
    [ $rep = 1 ] && t "$v c3-gi" CRT_B200_LIB=$lib timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
  done
done
cat $out
