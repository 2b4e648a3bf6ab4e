#!/bin/bash
# parity suite on the default build, A/B against the variants given, SAH triangle-cost sweep on the default build
tag=${1:-r2z}; shift
variants=${@:-"prev"}
mkdir -p gpurun_out
V=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_pytest_gpu.txt 2>&1; echo "suite exit $?" >> gpurun_out/${tag}_pytest_gpu.txt
tail -4 gpurun_out/${tag}_pytest_gpu.txt
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | grep -o "'gpu_ms': [0-9.]*")" >> $out; }
for rep in 1 2; do
  for v in default $variants; do
    lib=$V/libcrt_b200_$v.so; [ $v = default ] && lib=$PWD/chaos-ray-tracing-course-2025_b200/lib/libcrt_b200.so
    t "$v c4" CRT_B200_LIB=$lib timeout 200 python tools/render_frames.py c4 6 auto kd 0
    t "$v c4 rank0of8" CRT_B200_LIB=$lib CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 auto kd 0
    [ $rep = 1 ] && t "$v c5" CRT_B200_LIB=$lib timeout 200 python tools/render_frames.py c5 6 auto kd 0
    [ $rep = 1 ] && t "$v c3" CRT_B200_LIB=$lib timeout 200 python tools/render_frames.py c3 6 auto kd 0
    [ $rep = 1 ] && t "$v c1" CRT_B200_LIB=$lib timeout 200 python tools/render_frames.py c1 6 auto kd 0
    [ $rep = 1 ] && t "$v c3-gi" CRT_B200_LIB=$lib timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
  done
  for tri in 0.7 1.0 1.5 2.0 2.8; do
    t "sah tri=$tri c4" CRT_SAH_TRI=$tri timeout 200 python tools/render_frames.py c4 6 auto kd 0
    [ $rep = 1 ] && t "sah tri=$tri c4 rank0of8" CRT_SAH_TRI=$tri CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 auto kd 0
    [ $rep = 1 ] && t "sah tri=$tri c3-gi" CRT_SAH_TRI=$tri timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
  done
done
cat $out
