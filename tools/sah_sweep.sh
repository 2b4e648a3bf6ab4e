#!/bin/bash
# frame time and test counts of a workload for a few cost settings of the SAH walk-tree builder   usage: tools/sah_sweep.sh <workload> [pipeline]
w=${1:-c4}
p=${2:-auto}
for cfg in "1 2 2.8 0.85" "1 1 1.5 0.85" "1 0.5 1 0.85" "1 0.3 0.6 0.85" "1 3 5 0.85" "1 2 2.8 1.0" "1 1 1.5 1.0" "1 0.5 1 1.0"; do
  set -- $cfg
  echo -n "node=$1 leaf=$2 tri=$3 empty=$4: "
  CRT_SAH_NODE=$1 CRT_SAH_LEAF=$2 CRT_SAH_TRI=$3 CRT_SAH_EMPTY=$4 timeout 120 python tools/render_frames.py $w 3 $p kd 0 | tail -1 | cut -c1-60 | tr '\n' ' '
  CRT_SAH_NODE=$1 CRT_SAH_LEAF=$2 CRT_SAH_TRI=$3 CRT_SAH_EMPTY=$4 timeout 120 python tools/render_frames.py $w 1 $p kd 1 | tail -1 | grep -o "'node_tests.*"
done
