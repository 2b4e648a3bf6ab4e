#!/bin/bash
# quick check + profile of the current build: pytest -m gpu, C4 / C1 timings, ncu full capture of the C4 walk kernels
tag=${1:-prof}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_pytest_gpu.txt 2>&1; echo "suite exit $?" >> gpurun_out/${tag}_pytest_gpu.txt
tail -4 gpurun_out/${tag}_pytest_gpu.txt
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | grep -o "'gpu_ms': [0-9.]*")" >> $out; }
for rep in 1 2; do for w in c4 c1 c3 c5; do t "$w" timeout 200 python tools/render_frames.py $w 6 auto kd 0; done; done
cat $out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${tag}_launches_c4.csv python tools/render_frames.py c4 3 auto kd 0 > /dev/null 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"k_wf_trace|k_render_persistent" -c 3 -f -o gpurun_out/${tag}_trace_c4 \
  python tools/render_frames.py c4 1 auto kd 0 > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
