#!/usr/bin/env python
"""Writes the per-frame ray counts of the bench workloads (tests/golden/ray_counts.json): what bench.py's reference arm
divides by -- the reference renderer cannot count its own rays, and the arm must not load the product to count them.
Counted by the device's counting build (collect_stats); tests/test_ray_counts.py pins the table to the oracle port on
the CPU (c1, c5 frames, a c4 sample) and the GPU suite to the device (every workload).

    python tools/count_rays.py [out.json]      (needs a GPU)
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

import bench  # noqa: E402
import __graft_entry__ as entry  # noqa: E402

crt = entry.load_package()
from crt_b200 import scenes  # noqa: E402

out = {}
r = crt.Renderer(0)
for workload, (desc, W, H, depth, _) in bench.WORKLOADS.items():
    text, root = bench.scene_text(workload)
    flat = crt.Scene.load_json(text, root).flatten(reference_tree=False)
    r.upload(flat)
    params = crt.default_params(max_ray_depth=depth, collect_stats=1)
    entry_ = {"workload": desc, "width": W, "height": H, "max_ray_depth": depth}
    if workload == "c5":
        center = (np.array(list(flat.root_min)) + np.array(list(flat.root_max))) / 2
        orbit = scenes.orbit_cameras(list(flat.camera.position), list(flat.camera.rotation), center, 240)
        counts = []
        for pos, mat in orbit:
            cam = crt.Camera.from_buffer_copy(flat.camera)
            cam.position[:] = [float(v) for v in pos]
            cam.rotation[:] = [float(v) for v in mat]
            r.set_camera(cam)
            _, _, st = r.render(params, want_rgb=False, want_rgb8=False)
            counts.append(int(st["rays"]))
        entry_["rays_per_orbit_frame"] = counts
        entry_["rays_per_frame"] = int(round(sum(counts) / len(counts)))
    else:
        _, _, st = r.render(params, want_rgb=False, want_rgb8=False)
        entry_["rays_per_frame"] = int(st["rays"])
        entry_["rays_by_depth"] = [int(v) for v in st["rays_by_depth"] if v]
    out[workload] = entry_
    print(workload, entry_["rays_per_frame"], flush=True)
r.close()
path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "ray_counts.json")
json.dump(out, open(path, "w"), indent=1)
print("wrote", path)
