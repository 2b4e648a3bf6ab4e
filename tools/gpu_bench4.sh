#!/bin/bash
# bench lines of the shipped build (default workload with the CPU sample; C1, C3, C5 without)
tag=${1:-r3i}
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"; cut -c1-300 gpurun_out/${tag}_bench.json
for w in c1 c3 c5; do timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err; echo "bench $w exit $?"; cut -c1-250 gpurun_out/${tag}_bench_$w.json; done
