#!/bin/bash
# frame time of a workload for every tuned build under lib/variants/
w=${1:-c4}
for lib in chaos-ray-tracing-course-2025_b200/lib/variants/*.so; do
  tag=$(basename $lib .so | sed 's/libcrt_b200_//')
  CRT_B200_LIB=$PWD/$lib timeout 120 python tools/render_frames.py $w 3 wavefront kd 0 | tail -1 | sed "s/^/$tag: /" | cut -c1-80
done
