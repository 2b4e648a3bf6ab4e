#!/bin/bash
# step 29: parity suite on the default build (eager misses), then A/B against the build before and two shade triggers
tag=${1:-r3k}; shift
variants=$@
mkdir -p gpurun_out
V=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_pytest_gpu.txt 2>&1; echo "suite exit $?" >> gpurun_out/${tag}_pytest_gpu.txt
tail -3 gpurun_out/${tag}_pytest_gpu.txt
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | grep -o "'gpu_ms': [0-9.]*")" >> $out; }
for rep in 1 2; do
  for v in default $variants; do
    lib=$V/libcrt_b200_$v.so; [ $v = default ] && lib=$PWD/chaos-ray-tracing-course-2025_b200/lib/libcrt_b200.so
    t "$v c4" CRT_B200_LIB=$lib timeout 200 python tools/render_frames.py c4 6 auto kd 0
    t "$v c4 rank0of8" CRT_B200_LIB=$lib CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 auto kd 0
    [ $rep = 1 ] && [ $v != e16 ] && [ $v != e8 ] && t "$v c5 depth 8" CRT_B200_LIB=$lib CRT_DEPTH=8 timeout 200 python tools/render_frames.py c5 6 auto kd 0
  done
done
cat $out
