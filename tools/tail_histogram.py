#!/usr/bin/env python
"""Debug build only (make NVEXTRA=-DCRT_DEBUG_TAIL): rounds of the trace loop per ray, per bounce level."""
import ctypes as C
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench, __graft_entry__ as entry
crt = entry.load_package()
workload = sys.argv[1] if len(sys.argv) > 1 else "c4"
desc, W, H, depth, _ = bench.WORKLOADS[workload]
text, root = bench.scene_text(workload)
r = crt.Renderer(0); r.upload(crt.Scene.load_json(text, root))
p = crt.default_params(max_ray_depth=depth, pipeline=2)
buf = (C.c_ulonglong * 640)()
r.render(p, want_rgb=False, want_rgb8=False); crt.lib.crt_debug_tail(buf, 1)
_, _, st = r.render(p, want_rgb=False, want_rgb8=False); crt.lib.crt_debug_tail(buf, 1)
print("gpu_ms", st["gpu_ms"])
for lvl in range(depth + 1):
    row = buf[lvl * 40:(lvl + 1) * 40]
    n = sum(row[2:])
    print(f"level {lvl}: rays {n} max rounds {row[0]} mean {row[1] / max(n, 1):.1f} log2 histogram {[int(v) for v in row[2:20]]}"
          f"  | us until queue empty (max over warps) {row[30] / 1e3:.0f}, until last warp left {row[31] / 1e3:.0f}, mean warp lifetime {row[32] / max(row[33], 1) / 1e3:.0f}")
