#!/bin/bash
# round 2, pass j: GI kernel variants (out-of-line bounce direction, launch bounds, round length), inst_alg measurement
tag=${1:-r2j}
mkdir -p gpurun_out
V=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants
out=gpurun_out/${tag}_ab.txt
: > $out
t() { local label=$1; shift; echo "$label: $(env "$@" 2>&1 | tail -1 | cut -c1-70)" >> $out; }
for rep in 1 2; do
  t "default c3-gi" timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
  for v in gio gi7 gio7 gis4 gist20; do
    t "$v c3-gi" CRT_B200_LIB=$V/libcrt_b200_$v.so timeout 300 python tools/render_frames.py c3-gi 2 auto kd 0
  done
done
cat $out
timeout 1500 python tools/inst_alg.py gpurun_out/${tag}_inst_alg.json > gpurun_out/${tag}_inst_alg.log 2>&1; echo "inst_alg exit $?"; tail -12 gpurun_out/${tag}_inst_alg.log
for mode in 0 2 1; do
  echo "CRT_EARLY_COPY=$mode: $(CRT_EARLY_COPY=$mode timeout 300 python tools/e2e_probe.py c4 20 2>&1 | grep -E 'float frame' )" | tee -a gpurun_out/${tag}_early_copy.txt
  CRT_EARLY_COPY=$mode timeout 300 python - <<'PY' 2>&1 | tee -a gpurun_out/${tag}_early_copy.txt
import os, sys
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import torch, bench, __graft_entry__ as entry
crt = entry.load_package()
desc, W, H, depth, _ = bench.WORKLOADS["c4"]
text, root = bench.scene_text("c4")
r = crt.Renderer(0); r.upload(crt.Scene.load_json(text, root).flatten(reference_tree=False))
host = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
p = crt.default_params(max_ray_depth=depth)
ms = []
for k in range(13):
    _, _, st = r.render(p, want_rgb8=False, rgb_out=host.data_ptr())
    ms.append(st["gpu_ms"])
print("   kernel time inside crt_cuda_render (CUDA events around the frame's kernels), CRT_EARLY_COPY=%s: %.3f ms" % (os.environ.get("CRT_EARLY_COPY"), sum(ms[3:]) / 10))
PY
done
