#!/bin/bash
# quick GPU check used during kernel work: parity suite, then frame times of the bench workloads
tag=${1:-check}
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; tail -3 gpurun_out/${tag}_pytest.log
timeout 120 python tools/render_frames.py c4 3 wavefront kd 0 | tail -1
timeout 120 python tools/render_frames.py c4 1 wavefront kd 1
for w in c1 c3 c5; do timeout 100 python tools/render_frames.py $w 3 wavefront kd 0 | tail -1; done
