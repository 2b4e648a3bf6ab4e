#!/bin/bash
# round 2, pass e: L2 warm-up at the start of a frame, whole frame and one rank's share of a 2- / 8-rank frame
tag=${1:-r2e}
mkdir -p gpurun_out
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | cut -c1-70)" >> $out; }
for rep in 1 2; do
  for warm in 0 1; do
    t "warm=$warm c4 whole" CRT_WARM_L2=$warm timeout 200 python tools/render_frames.py c4 6 auto kd 0
    t "warm=$warm c4 rank0of2" CRT_WARM_L2=$warm CRT_RANK=0 CRT_WORLD=2 timeout 200 python tools/render_frames.py c4 6 auto kd 0
    t "warm=$warm c4 rank0of8" CRT_WARM_L2=$warm CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 auto kd 0
    t "warm=$warm c4 rank3of8 wavefront" CRT_WARM_L2=$warm CRT_RANK=3 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 wavefront kd 0
    t "warm=$warm c3 whole" CRT_WARM_L2=$warm timeout 200 python tools/render_frames.py c3 6 auto kd 0
    t "warm=$warm c5 whole" CRT_WARM_L2=$warm timeout 200 python tools/render_frames.py c5 6 auto kd 0
  done
done
cat $out
for warm in 0 1; do
  CRT_WARM_L2=$warm timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --sustain 0 > gpurun_out/${tag}_bench_warm$warm.json 2> gpurun_out/${tag}_bench_warm$warm.err; echo "bench warm=$warm exit $?"; cut -c1-330 gpurun_out/${tag}_bench_warm$warm.json
done
CRT_WARM_L2=1 CRT_RANK=0 CRT_WORLD=8 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_launches_c4_rank0of8_warm.csv \
    python tools/render_frames.py c4 3 auto kd 0 > gpurun_out/${tag}_ncu_rank0.log 2>&1; echo "ncu exit $?"
