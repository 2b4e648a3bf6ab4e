#!/bin/bash
# round 2, pass o (8 GPUs): bench.py as the driver launches it -- C4 at N = 8, 4, 2; 15-01/scene2 with GI and the C5
# orbit at N = 8; the hybrid hand-over level for one rank's share of an 8-rank frame
tag=${1:-r2t}
mkdir -p gpurun_out
run() { # n workload
  local n=$1 w=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $n --steps 20 --warmup 5 --workload $w "$@" \
    > gpurun_out/${tag}_bench_${w}_n$n.json 2> gpurun_out/${tag}_bench_${w}_n$n.err; echo "bench $w n$n exit $?"; cut -c1-260 gpurun_out/${tag}_bench_${w}_n$n.json; grep -v "resource_tracker\|warnings.warn\|OMP_NUM\|\*\*\*\*\|^$\|NCCL version" gpurun_out/${tag}_bench_${w}_n$n.err | tail -3
}
run 8 c4
run 4 c4
run 2 c4
run 8 c3-gi
run 8 c5
run 8 c3
for h in 1 2 3; do echo "hybrid level $h rank0of8: $(CRT_HYBRID_LEVEL=$h CRT_RANK=0 CRT_WORLD=8 python tools/render_frames.py c4 6 auto kd 0 2>&1 | tail -1 | cut -c1-70)"; done
timeout 600 python bench.py --workload c5 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_bench_c5_n1.json 2>/dev/null; cut -c1-260 gpurun_out/${tag}_bench_c5_n1.json
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_bench_c4_n1.json 2>/dev/null; cut -c1-260 gpurun_out/${tag}_bench_c4_n1.json
CRT_RANK=0 CRT_WORLD=8 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_launches_c4_rank0of8.csv python tools/render_frames.py c4 3 auto kd 0 > /dev/null 2>&1; echo "ncu exit $?"
