#!/bin/bash
# the seeded kernel now renders levels 1-8 of C4 (72 % of the frame): its round length, shade trigger and patience again;
# the SAH triangle cost again, three passes interleaved
tag=${1:-sweep3}; shift
variants=$@
mkdir -p gpurun_out
V=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | grep -o "'gpu_ms': [0-9.]*")" >> $out; }
for rep in 1 2 3; do
  for tri in 1.0 0.85 0.7 0.6; do t "sah tri=$tri c4" CRT_SAH_TRI=$tri timeout 200 python tools/render_frames.py c4 6 auto kd 0; done
  [ $rep = 3 ] && continue
  for v in $variants; do
    t "$v c4" CRT_B200_LIB=$V/libcrt_b200_$v.so timeout 200 python tools/render_frames.py c4 6 auto kd 0
  done
done
for tri in 1.0 0.7; do t "sah tri=$tri c3-gi" CRT_SAH_TRI=$tri timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0; t "sah tri=$tri c4 rank0of8" CRT_SAH_TRI=$tri CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 auto kd 0; done
cat $out
