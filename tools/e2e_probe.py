#!/usr/bin/env python
"""Where the end-to-end frame time goes: the device-only frame, and crt_cuda_render into a pinned host buffer (copy
queued behind the kernels, or started after the level-0 shade with CRT_EARLY_COPY=1).  usage: tools/e2e_probe.py [workload] [frames]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
import __graft_entry__ as entry  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c4"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 20
crt = entry.load_package()
desc, W, H, depth, _ = bench.WORKLOADS[workload]
text, root = bench.scene_text(workload)
scene = crt.Scene.load_json(text, root)
r = crt.Renderer(0)
r.upload(scene.flatten(reference_tree=False))
p = crt.default_params(max_ray_depth=depth)
host = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
host8 = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory()


def timed(label, fn):
    for _ in range(3):
        fn()
    t0 = time.perf_counter()
    for _ in range(frames):
        fn()
    ms = (time.perf_counter() - t0) / frames * 1e3
    print(f"{workload} {label}: {ms:.3f} ms/frame (CRT_EARLY_COPY={os.environ.get('CRT_EARLY_COPY', '0')})")


def device_only():
    r.render_device(p)
    r.synchronize()


timed("device only", device_only)
timed("float frame to host", lambda: r.render(p, want_rgb8=False, rgb_out=host.data_ptr()))
timed("quantised frame to host", lambda: r.render(p, want_rgb=False, rgb8_out=host8.data_ptr()))
timed("both to host", lambda: r.render(p, rgb_out=host.data_ptr(), rgb8_out=host8.data_ptr()))
r.close()
