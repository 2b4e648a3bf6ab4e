#!/bin/bash
# round 2, pass f (8 GPUs): bench.py as the driver launches it, N = 8: C4 in row bands, 15-01/scene2 with GI at 2160x2160
# in row bands, the C5 orbit sharded by frame; and N = 4 for C4.
tag=${1:-r2f}
mkdir -p gpurun_out
run() { # n workload extra...
  local n=$1 w=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $n --steps 20 --warmup 3 --workload $w "$@" \
    > gpurun_out/${tag}_bench_${w}_n$n.json 2> gpurun_out/${tag}_bench_${w}_n$n.err; echo "bench $w n$n exit $?"; cat gpurun_out/${tag}_bench_${w}_n$n.json; grep -v "resource_tracker\|warnings.warn\|OMP_NUM\|\*\*\*\*\|^$" gpurun_out/${tag}_bench_${w}_n$n.err | tail -3
}
run 8 c4
run 8 c3-gi
run 8 c5
run 4 c4
run 2 c4
