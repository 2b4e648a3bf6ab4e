#!/bin/bash
# round 2, pass b: A/B of the select-style lean walker against round 1's, launch-bound / round-length variants, parity
# suite, ncu of the lean kernels.  usage: tools/gpu_r2b.sh <tag>
tag=${1:-r2b}
mkdir -p gpurun_out
V=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | cut -c1-120)" >> $out; }
for rep in 1 2; do
  for w in c4 c1 c3 c5; do
    t "lean    $w" timeout 200 python tools/render_frames.py $w 5 auto kd 0
    t "classic $w" CRT_WALKER=classic CRT_B200_LIB=$V/libcrt_b200_classic.so timeout 200 python tools/render_frames.py $w 5 auto kd 0
  done
  for v in n9 n12 s6 s12 p8 n9p8 ps8; do
    t "$v      c4" CRT_B200_LIB=$V/libcrt_b200_$v.so timeout 200 python tools/render_frames.py c4 5 auto kd 0
  done
done
t "lean    c3-gi" timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
t "classic c3-gi" CRT_WALKER=classic CRT_B200_LIB=$V/libcrt_b200_classic.so timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
t "lean    c4 stats" timeout 200 python tools/render_frames.py c4 2 auto kd 1
cat $out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/${tag}_pytest.log
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"k_wf_trace|k_render_persistent" -c 3 -f -o gpurun_out/${tag}_trace_c4 \
  python tools/render_frames.py c4 1 auto kd 0 > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
