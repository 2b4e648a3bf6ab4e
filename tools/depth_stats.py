#!/usr/bin/env python
"""Per bounce level: rays, node steps, triangle tests and time of a workload (differences of renders at max_ray_depth 0..D)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench, __graft_entry__ as entry
workload = sys.argv[1] if len(sys.argv) > 1 else "c4"
D = int(sys.argv[2]) if len(sys.argv) > 2 else 3
crt = entry.load_package()
text, root = bench.scene_text(workload)
r = crt.Renderer(0)
r.upload(crt.Scene.load_json(text, root))
prev = None
for d in range(D + 1):
    _, _, st = r.render(crt.default_params(max_ray_depth=d, collect_stats=1, pipeline=2, traversal=1), want_rgb=False, want_rgb8=False)
    ms = min(r.render(crt.default_params(max_ray_depth=d, pipeline=2, traversal=1), want_rgb=False, want_rgb8=False)[2]["gpu_ms"] for _ in range(3))
    cur = (st["rays"], st["node_tests"], st["tri_tests"], ms)
    delta = cur if prev is None else tuple(a - b for a, b in zip(cur, prev))
    print(f"level {d}: rays {delta[0]:9d}  node steps/ray {delta[1] / max(1, delta[0]):7.2f}  tri tests/ray {delta[2] / max(1, delta[0]):6.2f}  +{delta[3]:.3f} ms")
    prev = cur
r.close()
