#!/bin/bash
# with the 58-instruction visit: hand-over level of the hybrid pipeline, batch triggers (variants) -- C4 and one rank's share
tag=${1:-sweep2}; shift
variants=$@
mkdir -p gpurun_out
V=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | grep -o "'gpu_ms': [0-9.]*")" >> $out; }
for rep in 1 2; do
  t "default c4" timeout 200 python tools/render_frames.py c4 6 auto kd 0
  [ $rep = 1 ] && t "default c4 rank0of8" CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 auto kd 0
  for h in 1 3 4; do t "hybrid level $h c4" CRT_HYBRID_LEVEL=$h timeout 200 python tools/render_frames.py c4 6 auto kd 0; done
  t "wavefront c4" timeout 200 python tools/render_frames.py c4 6 wavefront kd 0
  for v in $variants; do
    t "$v c4" CRT_B200_LIB=$V/libcrt_b200_$v.so timeout 200 python tools/render_frames.py c4 6 auto kd 0
    [ $rep = 1 ] && t "$v c4 rank0of8" CRT_B200_LIB=$V/libcrt_b200_$v.so CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 auto kd 0
  done
done
cat $out
