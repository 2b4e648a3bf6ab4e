#!/bin/bash
# ncu --set full of the GI persistent kernel on a 540x540 frame of 15-01/scene2 (the full frame would replay 0.2 s x 40)
tag=${1:-round}
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_render_persistent" -c 1 -f -o gpurun_out/${tag}_c3gi_540 \
  python tools/render_frames.py c3-gi 1 auto kd 0 540x540 > gpurun_out/${tag}_ncu_full_gi.log 2>&1; echo "ncu gi exit $?"
