#!/bin/bash
# frame time of a workload for every tuned build under lib/variants/   usage: tools/sweep.sh <workload> [pipeline]
w=${1:-c4}
p=${2:-wavefront}
for lib in chaos-ray-tracing-course-2025_b200/lib/variants/*.so; do
  tag=$(basename $lib .so | sed 's/libcrt_b200_//')
  CRT_B200_LIB=$PWD/$lib timeout 200 python tools/render_frames.py $w 3 $p kd 0 | tail -1 | sed "s/^/$tag: /" | cut -c1-80
done
