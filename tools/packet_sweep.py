#!/usr/bin/env python
"""Where warp packets (crt_trace_packet.cuh) beat the per-lane walk on the primary level: diffuse height fields of
decreasing triangle count at a fixed image size, frame time with CRT_PACKET=0 and 1 (the env var is read at upload).
usage: tools/packet_sweep.py [width height]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

crt = entry.load_package()
from crt_b200 import scenes  # noqa: E402

W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1080)
r = crt.Renderer(0)
for quads in (708, 500, 354, 250, 177, 125, 88, 44, 16):
    scene = crt.Scene.load_json(scenes.height_field_scene(quads, W, H, "diffuse"))
    flat = scene.flatten(reference_tree=False)
    line = f"{2 * quads * quads:8d} triangles, {W * H / (2 * quads * quads):8.1f} px/triangle:"
    for packet in ("0", "1"):
        os.environ["CRT_PACKET"] = packet
        r.upload(flat)
        p = crt.default_params(max_ray_depth=3)
        best = 1e9
        for _ in range(8):
            _, _, st = r.render(p, want_rgb=False, want_rgb8=False)
            best = min(best, st["gpu_ms"])
        line += f"  packet={packet} {best:7.3f} ms"
    print(line)
r.close()
