#!/bin/bash
# round 2, pass a: parity suite with the lean walker, A/B of walkers and launch-bound variants, ncu of the lean kernels,
# ray-count table.  usage: tools/gpu_r2a.sh <tag>
tag=${1:-r2a}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv > gpurun_out/${tag}_gpu.txt 2>&1
V=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants
out=gpurun_out/${tag}_ab.txt
: > $out
t() { # label, env..., workload
  local label=$1; shift
  echo "$label: $(env "$@" 2>&1 | tail -1 | cut -c1-200)" >> $out
}
for rep in 1 2; do
  for w in c4 c1 c3 c5; do
    t "lean    $w" timeout 200 python tools/render_frames.py $w 5 auto kd 0
    t "classic $w" CRT_WALKER=classic CRT_B200_LIB=$V/libcrt_b200_classic.so timeout 200 python tools/render_frames.py $w 5 auto kd 0
  done
  for v in n9 s12 p8; do
    t "$v      c4" CRT_B200_LIB=$V/libcrt_b200_$v.so timeout 200 python tools/render_frames.py c4 5 auto kd 0
  done
done
t "lean    c3-gi" timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
t "classic c3-gi" CRT_WALKER=classic CRT_B200_LIB=$V/libcrt_b200_classic.so timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
t "lean    c4 stats" timeout 200 python tools/render_frames.py c4 2 auto kd 1
t "lean    c4 wavefront" timeout 200 python tools/render_frames.py c4 5 wavefront kd 0
t "lean    c4 persistent" timeout 200 python tools/render_frames.py c4 5 persistent kd 0
cat $out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/${tag}_pytest.log
timeout 600 python tools/count_rays.py gpurun_out/${tag}_ray_counts.json > gpurun_out/${tag}_count_rays.log 2>&1; echo "count_rays exit $?"
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"; cat gpurun_out/${tag}_bench.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_c4.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"k_wf_trace|k_render_persistent" -c 3 -f -o gpurun_out/${tag}_trace_c4 \
  python tools/render_frames.py c4 1 auto kd 0 > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
ls -la gpurun_out | tail -12
