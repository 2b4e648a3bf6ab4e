#!/bin/bash
# eight ranks, the final build: one C4 frame in tiles, as the driver launches bench.py
tag=${1:-r3l}
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 20 --warmup 5 \
  > gpurun_out/${tag}_bench_c4_n8.json 2> gpurun_out/${tag}_bench_c4_n8.err; echo "bench c4 n8 exit $?"; cut -c1-300 gpurun_out/${tag}_bench_c4_n8.json
grep -v "resource_tracker\|warnings.warn\|OMP_NUM\|\*\*\*\*\|^$\|NCCL version" gpurun_out/${tag}_bench_c4_n8.err | tail -3
