#!/bin/bash
# round 2, pass g: the quad walker -- timing of every workload, parity suite, ncu of the C4 walk kernels
tag=${1:-r2g}
mkdir -p gpurun_out
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | cut -c1-100)" >> $out; }
for rep in 1 2; do
  for w in c4 c1 c3 c5; do t "quad $w" timeout 200 python tools/render_frames.py $w 6 auto kd 0; done
  t "quad c4 rank0of8" CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 auto kd 0
done
t "quad c3-gi" timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
t "quad c4 stats" timeout 200 python tools/render_frames.py c4 2 auto kd 1
cat $out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/${tag}_pytest.log
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"k_wf_trace|k_render_persistent" -c 3 -f -o gpurun_out/${tag}_trace_c4 \
  python tools/render_frames.py c4 1 auto kd 0 > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
