#!/usr/bin/env python
"""Render a few frames of a bench workload through the C ABI (no torch): the short command ncu wraps.
usage: python tools/render_frames.py <workload> [frames] [pipeline auto|mega|wavefront] [traversal kd|literal] [stats 0|1] [WxH]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import __graft_entry__ as entry  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c4"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 3
pipeline = sys.argv[3] if len(sys.argv) > 3 else "auto"
traversal = sys.argv[4] if len(sys.argv) > 4 else "kd"
stats = int(sys.argv[5]) if len(sys.argv) > 5 else 0
crt = entry.load_package()
desc, W, H, depth, _ = bench.WORKLOADS[workload]
depth = int(os.environ.get("CRT_DEPTH", depth))
size = tuple(int(v) for v in sys.argv[6].split("x")) if len(sys.argv) > 6 else None  # e.g. 540x540: a smaller frame for ncu
text, root = bench.scene_text(workload, size)
scene = crt.Scene.load_json(text, root)
r = crt.Renderer(0)
flat = r.upload(scene.flatten(reference_tree=traversal == "literal"))
# CRT_RANK / CRT_WORLD: render only the row bands rank r of an N-rank job owns (what each GPU of a tile-sharded frame does)
part = dict(rank=int(os.environ["CRT_RANK"]), world=int(os.environ["CRT_WORLD"]), band_rows=int(os.environ.get("CRT_BAND", "8"))) if os.environ.get("CRT_WORLD") else {}
p = crt.default_params(max_ray_depth=depth, collect_stats=stats, **part,
                       pipeline={"auto": 0, "mega": 1, "wavefront": 2, "persistent": 3, "hybrid": 4}[pipeline], traversal={"kd": 1, "literal": 0}[traversal])
flush = None
if os.environ.get("CRT_FLUSH") == "1":  # cold L2 before every frame, as bench.py's timed steps have it
    import torch
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
for i in range(frames):
    if flush is not None:
        flush.fill_(i & 0xFF)
        torch.cuda.synchronize()
    _, _, st = r.render(p, want_rgb=False, want_rgb8=False)
    print(workload, pipeline, traversal, "frame", i, {k: v for k, v in st.items() if v and k != "pad"})
r.close()
