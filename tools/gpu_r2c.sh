#!/bin/bash
# round 2, pass c: memory-side variants of the lean kernels (8-byte nodes, streamed frame / queue traffic, prefetch of
# pushed nodes), the new bench.py at N = 1.   usage: tools/gpu_r2c.sh <tag>
tag=${1:-r2c}
mkdir -p gpurun_out
V=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | cut -c1-70)" >> $out; }
for rep in 1 2 3; do
  for v in base node8 stream pref sn sp all3; do
    t "$v c4" CRT_B200_LIB=$V/libcrt_b200_$v.so timeout 200 python tools/render_frames.py c4 6 auto kd 0
  done
done
for v in base node8 stream sn all3; do
  for w in c1 c3 c5; do
    t "$v $w" CRT_B200_LIB=$V/libcrt_b200_$v.so timeout 200 python tools/render_frames.py $w 6 auto kd 0
  done
  t "$v c3-gi" CRT_B200_LIB=$V/libcrt_b200_$v.so timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
done
cat $out
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"; cat gpurun_out/${tag}_bench.json; tail -5 gpurun_out/${tag}_bench.err
timeout 600 python bench.py --workload c5 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_c5.json 2> gpurun_out/${tag}_bench_c5.err; echo "bench c5 exit $?"; cat gpurun_out/${tag}_bench_c5.json; tail -5 gpurun_out/${tag}_bench_c5.err
