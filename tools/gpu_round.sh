#!/bin/bash
# One GPU-box pass: parity suite, smoke, bench (both arms), then the ncu launch list and one --set full capture of the
# trace kernel.  Everything lands in gpurun_out/<tag>_*.  usage: tools/gpu_round.sh <tag> [skip-ncu]
tag=${1:-round}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv > gpurun_out/${tag}_gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${tag}_smoke.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"; cat gpurun_out/${tag}_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "ref exit $?"; cat gpurun_out/${tag}_bench_ref.json
for w in c1 c3 c5; do timeout 200 python tools/render_frames.py $w 3 auto kd 1 2>&1 | tail -1; done | tee gpurun_out/${tag}_other_workloads.txt
timeout 300 python tools/render_frames.py c3-gi 3 auto kd 1 2>&1 | tail -1 | tee -a gpurun_out/${tag}_other_workloads.txt
[ "$2" = "skip-ncu" ] && exit 0
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_c4.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"k_wf_trace|k_render_persistent" -c 3 -f -o gpurun_out/${tag}_trace_c4 \
  python tools/render_frames.py c4 1 auto kd 0 > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
ls -la gpurun_out | tail -20
