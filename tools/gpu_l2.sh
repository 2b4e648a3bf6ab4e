#!/bin/bash
# step 26: L2 set-aside + access policy window over the walk's working set ($CRT_L2_PERSIST_MB, $CRT_L2_WINDOW, $CRT_L2_MISS)
tag=${1:-l2}
mkdir -p gpurun_out
out=gpurun_out/${tag}_l2_policy.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | grep -o "'gpu_ms': [0-9.]*")" >> $out; }
CRT_L2_VERBOSE=1 CRT_L2_PERSIST_MB=64 timeout 200 python tools/render_frames.py c4 1 auto kd 0 2>&1 | grep "crt: L2" >> $out
for rep in 1 2; do
  for cfg in "0 all n" "32 all n" "64 all n" "96 all n" "128 all n" "40 tree n" "64 tree n" "64 all s" "64 tree s" "96 all s"; do
    set -- $cfg
    for flush in 0 1; do
      t "mb=$1 window=$2 miss=$3 flush=$flush c4" CRT_L2_PERSIST_MB=$1 CRT_L2_WINDOW=$2 CRT_L2_MISS=$3 CRT_FLUSH=$flush timeout 200 python tools/render_frames.py c4 6 auto kd 0
    done
    [ $rep = 1 ] && t "mb=$1 window=$2 miss=$3 c4 rank0of8" CRT_L2_PERSIST_MB=$1 CRT_L2_WINDOW=$2 CRT_L2_MISS=$3 CRT_RANK=0 CRT_WORLD=8 timeout 200 python tools/render_frames.py c4 6 auto kd 0
  done
done
cat $out
