#!/usr/bin/env python
"""SURVEY §8(d) "issue" roofline: inst_alg = the walk kernel's own NON-DIVERGENT warp instructions per node visit, per
triangle test and per ray, measured with ncu (smsp__inst_executed.sum) on FULLY COHERENT ray batches ($CRT_COHERENT=1:
the 32 pixels of a patch all take lane 0's ray, so every lane of a warp is at the same node at the same time).

For each sample (a primary-only render of a workload) two runs of the same frame:
  * a counting build in-process  -> N_node, N_tri (thread-level counts: 32 x the warp-level ones here), N_ray
  * the timed build under ncu    -> warp instructions of the k_wf_trace_lean launch
and then least squares for  I_warp = a * N_node/32 + b * N_tri/32 + c * N_ray/32.
Writes the fit into profiles/traffic.json (workload entries get "inst_alg") and prints the samples.

    python tools/inst_alg.py [out.json]      (needs a GPU and ncu)
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

os.environ["CRT_COHERENT"] = "1"
os.environ["CRT_PACKET"] = "0"  # the per-lane walk kernel also where the warp-packet kernel would render the primary level
import bench  # noqa: E402
import __graft_entry__ as entry  # noqa: E402

crt = entry.load_package()
SAMPLES = [("c4-diffuse", None), ("c4-diffuse", "1920x1080"), ("c1", None), ("c3", None), ("c5", None), ("c1", "3840x2160")]


def counted(workload, size):
    text, root = bench.scene_text(workload, tuple(int(v) for v in size.split("x")) if size else None)
    flat = crt.Scene.load_json(text, root).flatten(reference_tree=False)
    r = crt.Renderer(0)
    r.upload(flat)
    _, _, st = r.render(crt.default_params(max_ray_depth=0, collect_stats=1, pipeline=crt.PIPELINE_WAVEFRONT), want_rgb=False, want_rgb8=False)
    r.close()
    return st


def warp_instructions(workload, size):
    cmd = ["ncu", "--metrics", "smsp__inst_executed.sum,smsp__thread_inst_executed.sum", "--clock-control", "none", "-k", "regex:k_wf_trace_lean", "-c", "1", "--csv",
           sys.executable, os.path.join(ROOT, "tools", "render_frames.py"), workload, "1", "wavefront", "kd", "0"] + ([size] if size else [])
    out = subprocess.run(cmd, env=dict(os.environ, CRT_DEPTH="0", CRT_PACKET="0"), capture_output=True, text=True, timeout=900).stdout
    rows = [r for r in csv.reader(io.StringIO(out)) if len(r) > 10 and r[0].isdigit()]
    vals = {r[-3]: float(r[-1].replace(",", "")) for r in rows}
    return vals["smsp__inst_executed.sum"], vals["smsp__thread_inst_executed.sum"]


samples = []
for workload, size in SAMPLES:
    st = counted(workload, size)
    warp, thread = warp_instructions(workload, size)
    # level-0 walks only: entered rays, their node visits (one "visit" per ray is the root-box test of the ray
    # generation, not a node of the walk) and triangle tests
    n_node = st["node_tests"] - st["rays"]
    samples.append({"workload": workload, "size": size, "rays_entered": st["entered_rays"], "node_visits": n_node, "tri_tests": st["tri_tests"],
                    "warp_instructions": warp, "threads_per_instruction": thread / warp})
    print(samples[-1], flush=True)
A = np.array([[s["node_visits"] / 32.0, s["tri_tests"] / 32.0, s["rays_entered"] / 32.0] for s in samples])
y = np.array([s["warp_instructions"] for s in samples])
coef, res, rank, _ = np.linalg.lstsq(A, y, rcond=None)
if coef[2] < 0:  # rays and node visits are nearly collinear over these samples: fold the per-ray share into the visits
    coef2, *_ = np.linalg.lstsq(A[:, :2], y, rcond=None)
    coef = np.array([coef2[0], coef2[1], 0.0])
fit = A @ coef
out = {"per_node_visit": float(coef[0]), "per_triangle_test": float(coef[1]), "per_ray": float(coef[2]),
       "fit_relative_error_max": float(np.max(np.abs(fit - y) / y)), "samples": samples,
       "method": "ncu smsp__inst_executed.sum of k_wf_trace_lean on fully coherent batches (CRT_COHERENT=1, every lane of a warp walks the same ray), "
                 "least squares over the samples (tools/inst_alg.py); warp instructions per WARP-level visit / test / ray"}
print(json.dumps({k: v for k, v in out.items() if k != "samples"}, indent=1))
path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "inst_alg.json")
json.dump(out, open(path, "w"), indent=1)
