#!/bin/bash
# A/B of the default library against one tuned build (lib/variants/libcrt_b200_<variant>.so) on the bench workloads,
# then the parity suite on the default library.   usage: tools/gpu_ab_round.sh <tag> <variant>
tag=${1:-ab}
variant=${2:-node16}
mkdir -p gpurun_out
vlib=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants/libcrt_b200_$variant.so
out=gpurun_out/${tag}_ab.txt
: > $out
for w in c4 c1 c3 c5; do
  for rep in 1 2; do
    echo "default $w: $(timeout 200 python tools/render_frames.py $w 4 auto kd 0 2>&1 | tail -1 | cut -c1-160)" >> $out
    echo "$variant $w: $(CRT_B200_LIB=$vlib timeout 200 python tools/render_frames.py $w 4 auto kd 0 2>&1 | tail -1 | cut -c1-160)" >> $out
  done
done
echo "default c3-gi: $(timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0 2>&1 | tail -1 | cut -c1-160)" >> $out
echo "$variant c3-gi: $(CRT_B200_LIB=$vlib timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0 2>&1 | tail -1 | cut -c1-160)" >> $out
cat $out
