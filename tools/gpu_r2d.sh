#!/bin/bash
# round 2, pass d (2 GPUs): the new bench at N = 2 (C4 in row bands, C5 orbit frames), the new GPU tests, and the
# per-kernel timeline of ONE rank of an 8-rank tile-sharded C4 frame (rank r of 8 played on one GPU under ncu).
tag=${1:-r2d}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_render_frame_gpu.py tests/test_ray_counts.py tests/test_pycrt.py tests/test_fullsize_gpu.py -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/${tag}_pytest.log
for w in c4 c5; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 --workload $w \
    > gpurun_out/${tag}_bench_${w}_n2.json 2> gpurun_out/${tag}_bench_${w}_n2.err; echo "bench $w n2 exit $?"; cat gpurun_out/${tag}_bench_${w}_n2.json; tail -3 gpurun_out/${tag}_bench_${w}_n2.err
done
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_c4_n1.json 2> gpurun_out/${tag}_bench_c4_n1.err; echo "bench n1 exit $?"; cat gpurun_out/${tag}_bench_c4_n1.json
for r in 0 3; do
  CRT_RANK=$r CRT_WORLD=8 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_launches_c4_rank${r}of8.csv \
    python tools/render_frames.py c4 3 auto kd 0 > gpurun_out/${tag}_ncu_rank${r}.log 2>&1; echo "ncu rank $r exit $?"
done
for r in 0 3; do echo "rank $r of 8: $(CRT_RANK=$r CRT_WORLD=8 python tools/render_frames.py c4 6 auto kd 0 2>&1 | tail -1 | cut -c1-100)"; done
echo "rank 0 of 2: $(CRT_RANK=0 CRT_WORLD=2 python tools/render_frames.py c4 6 auto kd 0 2>&1 | tail -1 | cut -c1-100)"
