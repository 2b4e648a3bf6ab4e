#!/bin/bash
# two ranks, the final build: one C4 frame in tiles, as the driver launches bench.py
tag=${1:-r3j}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 \
  > gpurun_out/${tag}_bench_c4_n2.json 2> gpurun_out/${tag}_bench_c4_n2.err; echo "bench c4 n2 exit $?"; cut -c1-300 gpurun_out/${tag}_bench_c4_n2.json
grep -v "resource_tracker\|warnings.warn\|OMP_NUM\|\*\*\*\*\|^$\|NCCL version" gpurun_out/${tag}_bench_c4_n2.err | tail -3
