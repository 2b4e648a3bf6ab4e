#!/bin/bash
# parameter sweep of the lean kernels on C4 (variants under lib/variants, hand-over level of the hybrid pipeline)
tag=${1:-sweep}
mkdir -p gpurun_out
V=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | cut -c1-60)" >> $out; }
for rep in 1 2; do
  t "default c4" timeout 200 python tools/render_frames.py c4 6 auto kd 0
  for v in r28 r20 t24 t10 ps4 ps8 st20 st6 pa4 pa16 w4 w10; do
    t "$v c4" CRT_B200_LIB=$V/libcrt_b200_$v.so timeout 200 python tools/render_frames.py c4 6 auto kd 0
  done
  for h in 1 3 4 8; do t "hybrid level $h c4" CRT_HYBRID_LEVEL=$h timeout 200 python tools/render_frames.py c4 6 auto kd 0; done
  t "wavefront c4" timeout 200 python tools/render_frames.py c4 6 wavefront kd 0
done
cat $out
