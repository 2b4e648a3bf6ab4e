#!/bin/bash
w=${1:-c4}
for cfg in "0 1 2 2.8" "1 1 2 2.8" "1 1 2 2.0" "1 1 2 1.5" "1 1 2 1.0" "1 1 6 2.0" "1 2 4 2.0" "1 1 2 0.7"; do
  set -- $cfg
  echo -n "collapse=$1 node=$2 leaf=$3 tri=$4: "
  CRT_TREE_COLLAPSE=$1 CRT_SAH_NODE=$2 CRT_SAH_LEAF=$3 CRT_SAH_TRI=$4 timeout 120 python tools/render_frames.py $w 3 wavefront kd 0 | tail -1 | cut -c1-70
  CRT_TREE_COLLAPSE=$1 CRT_SAH_NODE=$2 CRT_SAH_LEAF=$3 CRT_SAH_TRI=$4 timeout 120 python tools/render_frames.py $w 1 wavefront kd 1 | tail -1 | grep -o "'node_tests.*"
done
