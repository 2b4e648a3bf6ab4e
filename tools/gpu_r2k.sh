#!/bin/bash
# round 2, pass k: GI kernel at 7 blocks / 72 registers and variants around it; inst_alg; ncu of the GI kernel at 540x540
tag=${1:-r2k}
mkdir -p gpurun_out
V=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | cut -c1-70)" >> $out; }
for rep in 1 2; do
  t "default(g7) c3-gi" timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
  for v in g6 g7st20 g7st24 g7s3 g7s1 g7pa8 g7pa32 g7o; do
    t "$v c3-gi" CRT_B200_LIB=$V/libcrt_b200_$v.so timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
  done
done
cat $out
timeout 1500 python tools/inst_alg.py gpurun_out/${tag}_inst_alg.json > gpurun_out/${tag}_inst_alg.log 2>&1; echo "inst_alg exit $?"; tail -14 gpurun_out/${tag}_inst_alg.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_render_persistent" -c 1 -f -o gpurun_out/${tag}_c3gi_540 \
  python tools/render_frames.py c3-gi 1 auto kd 0 540x540 > gpurun_out/${tag}_ncu_full_gi.log 2>&1; echo "ncu gi exit $?"
