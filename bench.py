#!/usr/bin/env python
"""bench.py -- the headline benchmark of the render path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c4|c4-diffuse|c1|c3|c5]

A step is one frame of the workload.  Default workload (BASELINE.json configs[3], the one the metric's "1/2/4/8 B200"
is quoted on): the seeded 1 002 528-triangle height field, 3840x2160, max_ray_depth 8, one point light, reflective.
N > 1 (torchrun, one rank per GPU): the frame is cut into interleaved row bands, every rank renders its bands from a
replicated scene, and the quantised bands are gathered to rank 0 over NCCL inside the timed step ("strong" scaling:
total work is fixed).  Prints ONE JSON line on rank 0.

  value       Mrays/s, whole job, scene resident in HBM, device-timed (CUDA events on the launch stream), max over ranks
  e2e         same metric through the C ABI with HOST buffers: per step the camera/params go host->device with the
              launch and the float framebuffer comes back into pinned host memory (crt_cuda_render)
  roofline    dominant kernel against the measured HBM peak (MEASURED_PEAKS.json) using the ALGORITHMIC bytes
              32*N_node(16 for k-d nodes)+48*N_tri+16 per ray, with N_node/N_tri counted by the same kernel
  cpu_baseline the unmodified reference (oracle/_ref) on this box's host cores on a bounded sample (rank 0, N=1 only)

--impl reference times the reference's own CPU renderer (oracle/_ref/ref_harness -> crt::render_image) on a bounded
sample of the same workload; no kernel of ours runs on that arm.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

_JSON_OUT = None  # N > 1: a duplicate of the original stdout (see run_ours)

WORKLOADS = {
    # name: (description, width, height, max_ray_depth, sample size for the CPU arm)
    "c4": ("synthetic 1002528-triangle height field, reflective, 3840x2160, max_ray_depth 8, 1 point light", 3840, 2160, 8, (1920, 1080)),
    "c4-diffuse": ("synthetic 1002528-triangle height field, diffuse, 3840x2160, max_ray_depth 8, 1 point light", 3840, 2160, 8, (1920, 1080)),
    "c1": ("scenes/14-01-acceleration-tree/scene1 (== 13-01-optimizations/scene0), 1920x1080, defaults", 1920, 1080, 3, (1920, 1080)),
    "c3": ("scenes/15-01-conclusion/scene1 at 3840x2160, defaults", 3840, 2160, 3, (1920, 1080)),
    "c3-gi": ("scenes/15-01-conclusion/scene2 (GI on, refractive + reflective) at 2160x2160, defaults", 2160, 2160, 3, (360, 360)),
    "c5": ("scenes/11-01-refractive/scene8 1920x1080, one orbit frame per step (240-frame orbit, frame k -> rank k % N)", 1920, 1080, 3, (1920, 1080)),
}


def scene_text(workload: str, size=None) -> tuple[str, str]:
    """(json text, asset root) of the workload, optionally re-sized (CPU sample)."""
    import oracle_tools as ot  # only used here for the bundled-scene fixtures' location
    if workload.startswith("c4"):
        import __graft_entry__ as entry
        entry.load_package()
        from crt_b200 import scenes
        w, h = size or WORKLOADS[workload][1:3]
        return scenes.height_field_scene(708, w, h, "diffuse" if workload.endswith("diffuse") else "reflective"), ""
    rel = {"c1": "14-01-acceleration-tree/scene1.crtscene", "c3": "15-01-conclusion/scene1.crtscene", "c3-gi": "15-01-conclusion/scene2.crtscene", "c5": "11-01-refractive/scene8.crtscene"}[workload]
    w, h = size or WORKLOADS[workload][1:3]
    return ot.scene_variant(rel, w, h), os.path.dirname(ot.scene_file(rel))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.rows, self.proc, self.nvml = [], None, None
        # NVML in a thread (one sample every ~2 ms: a 20-step timed region lasts 60 ms); nvidia-smi -lms 20 otherwise
        try:
            import pynvml
            pynvml.nvmlInit()
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            pynvml.nvmlDeviceGetClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
            self.nvml, self.running = pynvml, True
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _poll(self):
        n = self.nvml
        bits = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))  # nvmlClocksEventReasons
        try:
            mx = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
        except Exception:
            mx = None
        while self.running:
            try:
                sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                try:
                    mask = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                except Exception:
                    mask = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                self.rows.append((time.perf_counter(), [str(sm), str(mx), "", *["Active" if mask & b else "Not Active" for _, b in bits]]))
            except Exception:
                pass
            time.sleep(0.002)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self, t0: float, t1: float) -> dict:
        if self.nvml:
            self.running = False
            self.thread.join(timeout=1.0)
        elif not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            time.sleep(0.15)
            self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1] or [r for _, r in self.rows[-3:]]
        sm, mx, reasons = [], [], set()
        for r in rows:
            try:
                sm.append(float(r[0])), mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm), "source": "nvml" if self.nvml else "nvidia-smi"}


def reference_sample(workload: str, steps: int, warmup: int) -> dict:
    """the reference's own crt::render_image on this box's host cores, on a bounded sample of the workload"""
    import __graft_entry__ as entry
    import oracle_tools as ot
    crt = entry.load_package()
    desc, W, H, depth, sample = WORKLOADS[workload]
    text, root = scene_text(workload, sample)
    if not ot.have_reference_build():
        # the oracle port stands in (kind "port") when oracle/_ref did not travel
        scene = crt.Scene.load_json(text, root)
        flat = scene.flatten()
        params = crt.default_params(max_ray_depth=depth, traversal=crt.TRAVERSAL_LITERAL)
        times = []
        for _ in range(warmup + steps):
            t0 = time.perf_counter()
            out = ot.port_render(flat, params, primary=False)
            times.append(time.perf_counter() - t0)
        rays, kind, cores = out["stats"]["rays"], "port", os.cpu_count()
        times = times[warmup:]
    else:
        with tempfile.TemporaryDirectory() as tmp:
            path = os.path.join(tmp, "sample.crtscene")
            open(path, "w").write(text)
            out = ot.run_harness(path, asset_root=root or None, max_depth=depth, repeat=warmup + steps)
        assert out.get("loaded"), out
        times, kind, cores = out["render_s"][warmup:], "reference", out["threads"]
        # the reference cannot count rays; the oracle port (bit-identical recursion) counts them for the same sample
        scene = crt.Scene.load_json(text, root)
        rays = ot.port_render(scene.flatten(), crt.default_params(max_ray_depth=depth, traversal=crt.TRAVERSAL_LITERAL), primary=False)["stats"]["rays"]
    sec = statistics.mean(times)
    return {"value": rays / sec / 1e6, "unit": "Mrays/s", "cores": cores, "kind": kind,
            "sample": f"{sample[0]}x{sample[1]} render of the same scene and camera ({rays} rays/frame), mean of {len(times)} frames, "
                      f"{sec * 1e3:.1f} ms/frame", "ms_per_step": sec * 1e3, "rays_per_step": rays}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ref = reference_sample(args.workload, args.steps, args.warmup)
    desc = WORKLOADS[args.workload][0]
    line = {"impl": "reference", "metric": "Mrays/s (primary+secondary)", "value": ref["value"], "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ref["ms_per_step"], "higher_is_better": True, "scaling": "weak" if args.partition == "frames" else "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "sample": ref["sample"]},
            "cpu_baseline": {"value": ref["value"], "unit": "Mrays/s", "cores": ref["cores"], "kind": ref["kind"], "sample": ref["sample"]},
            "e2e": {"value": ref["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist

    import __graft_entry__ as entry
    crt = entry.load_package()
    from crt_b200 import distributed as D
    from crt_b200 import scenes

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path is the CUDA library, there is no CPU fallback")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # NCCL's version / warning lines stay off stdout: ONE JSON line there
        # ... and since NCCL still writes its version line to file descriptor 1 on some boxes (r1zz, N = 2), the
        # descriptor itself points at stderr from here on; the JSON line goes to a duplicate of the original stdout
        global _JSON_OUT
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=device)

    desc, W, H, depth, _ = WORKLOADS[args.workload]
    t0 = time.perf_counter()
    text, root = scene_text(args.workload)
    scene = crt.Scene.load_json(text, root)
    flat = scene.flatten(reference_tree=False)  # the timed path walks the SAH tree only (TRAVERSAL_KD)
    t_load = time.perf_counter() - t0
    renderer = crt.Renderer(local)
    stream = torch.cuda.current_stream(device)
    renderer.set_stream(stream.cuda_stream)
    t0 = time.perf_counter()
    renderer.upload(flat)
    torch.cuda.synchronize(device)
    t_upload = time.perf_counter() - t0

    orbit = None
    if args.workload == "c5":
        a = flat.arrays()
        center = (np.array(list(flat.root_min)) + np.array(list(flat.root_max))) / 2
        orbit = scenes.orbit_cameras(list(flat.camera.position), list(flat.camera.rotation), center, 240)
    # N > 1: whole frames per rank with no exchange ("frames": animation-style, frame k -> rank k % N; for the still
    # workloads every frame uses the workload's camera, so each GPU does exactly the N = 1 work -> weak scaling), or one
    # frame cut into interleaved row bands that are pushed to rank 0 ("tiles": strong scaling; also measured below)
    frame_parallel = orbit is not None or (world > 1 and args.partition == "frames")
    band_rows = 8
    pipeline = {"auto": crt.PIPELINE_AUTO, "mega": crt.PIPELINE_MEGAKERNEL, "wavefront": crt.PIPELINE_WAVEFRONT, "persistent": crt.PIPELINE_PERSISTENT,
                "hybrid": crt.PIPELINE_HYBRID}[args.pipeline]
    common = dict(max_ray_depth=depth, traversal=crt.TRAVERSAL_KD, pipeline=pipeline)
    part = dict(rank=0, world=1) if frame_parallel else dict(rank=rank, world=world, band_rows=band_rows)
    params = crt.default_params(**common, **part)

    def camera_for(step):
        cam = crt.Camera.from_buffer_copy(flat.camera)
        if orbit is not None:
            pos, mat = orbit[(step * world + rank) % len(orbit)]
            cam.position[:] = [float(v) for v in pos]
            cam.rotation[:] = [float(v) for v in mat]
        return cam

    # --- ray / test counts of the timed configuration (counting build of the same kernels, outside the timed region)
    counted = {"rays": 0, "node_tests": 0, "tri_tests": 0}
    n_count_frames = min(8, args.steps) if orbit is not None else 1
    rays_per_step = []
    for s in range(n_count_frames):
        renderer.set_camera(camera_for(s))
        _, _, st = renderer.render(crt.default_params(collect_stats=1, **common, **part), want_rgb=False, want_rgb8=False)
        rays_per_step.append(st["rays"])
        for k in counted:
            counted[k] += st[k]
    kernel_launches = st["kernel_launches"]
    local_rays = torch.tensor([float(np.mean(rays_per_step))], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(local_rays)
    job_rays_per_step = float(local_rays.item())  # all ranks together: one frame (stills) or N frames (animation)

    gather, peer = None, None
    if world > 1 and not frame_parallel:
        if args.gather == "peer":
            peer = D.PeerFrame(renderer, W, H, device)  # bands are pushed into rank 0's frame over NVLink (crt_push.cuh)
        else:
            gather = D.BandGather(W, H, band_rows, 3, device)  # NCCL gather + un-interleave on rank 0
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)  # > 126 MB L2

    def device_step(step):
        renderer.set_camera(camera_for(step))
        renderer.render_device(params)
        if peer is not None:
            peer.push()
        elif gather is not None:
            _, rgb8_ptr, w, rows = renderer.device_framebuffer()
            gather.gather(D.wrap_device_bytes(rgb8_ptr, w * rows * 3, device))

    for s in range(args.warmup):
        device_step(s)
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    begin = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    end = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    kernel_ms = []
    wall0 = time.perf_counter()
    for s in range(args.steps):
        flush.fill_(s & 0xFF)  # L2 flush between timed iterations (outside the event bracket)
        begin[s].record(stream)
        device_step(args.warmup + s)
        end[s].record(stream)
        if args.steps <= 64:
            kernel_ms.append(renderer.synchronize()["gpu_ms"])
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()
    wall1 = time.perf_counter()
    step_ms = [b.elapsed_time(e) for b, e in zip(begin, end)]
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    clocks = sampler.stop(wall0, wall1) if sampler else None

    # --- e2e: the C-ABI call with host buffers (pinned), copies inside the timed region
    rows_local = len(crt.Renderer.owned_rows(H, params.rank, params.world, params.band_rows))
    host_rgb = torch.empty((rows_local, W, 3), dtype=torch.float32).pin_memory()
    host_ptr = host_rgb.data_ptr()
    e2e_steps = args.steps

    def e2e_step(step):
        renderer.set_camera(camera_for(step))
        renderer.render(params, want_rgb8=False, rgb_out=host_ptr)

    for s in range(min(3, args.warmup)):
        e2e_step(s)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for s in range(e2e_steps):
        e2e_step(args.warmup + s)
    t_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    t_e2e = float(t_e2e.item())

    # --- N > 1, still workloads: the tile-sharded single frame as well (strong scaling, exchange inside the timed step)
    tiles = None
    if world > 1 and orbit is None:
        t_params = crt.default_params(**common, rank=rank, world=world, band_rows=band_rows)
        t_peer = D.PeerFrame(renderer, W, H, device) if args.gather == "peer" else None
        t_gather = None if t_peer else D.BandGather(W, H, band_rows, 3, device)

        def tile_step():
            renderer.render_device(t_params)
            if t_peer is not None:
                t_peer.push()
            else:
                _, rgb8_ptr, w, rows = renderer.device_framebuffer()
                t_gather.gather(D.wrap_device_bytes(rgb8_ptr, w * rows * 3, device))

        renderer.set_camera(camera_for(0))
        for _ in range(3):
            tile_step()
        torch.cuda.synchronize(device)
        dist.barrier()
        t_steps = min(args.steps, 20)
        tb = [torch.cuda.Event(enable_timing=True) for _ in range(t_steps)]
        te = [torch.cuda.Event(enable_timing=True) for _ in range(t_steps)]
        for s in range(t_steps):
            flush.fill_(s & 0xFF)
            tb[s].record(stream)
            tile_step()
            te[s].record(stream)
        torch.cuda.synchronize(device)
        dist.barrier()
        t_ms = torch.tensor([sum(a.elapsed_time(c) for a, c in zip(tb, te)) / t_steps], dtype=torch.float64, device=device)
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
        frame_rays = job_rays_per_step / world if frame_parallel else job_rays_per_step
        tiles = {"ms_per_frame": float(t_ms.item()), "value": frame_rays / (float(t_ms.item()) * 1e-3) / 1e6, "unit": "Mrays/s", "scaling": "strong",
                 "partition": f"one frame, interleaved row bands of {band_rows}, " + ("pushed into rank 0's frame by peer stores over NVLink (crt_push.cuh) + a 1-element all-reduce as the barrier"
                                                                                       if args.gather == "peer" else "NCCL gather of u8 rgb to rank 0")}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    ms_per_step = total_ms / args.steps
    value = job_rays_per_step / (ms_per_step * 1e-3) / 1e6
    e2e_value = job_rays_per_step * e2e_steps / t_e2e / 1e6

    # roofline of the dominant kernel (the render kernel(s) of one frame on this rank)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    else:
        peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
    n = n_count_frames
    node_bytes = crt.lib.crt_cuda_walk_node_bytes()  # 8: the device records of this build (16 with -DCRT_NODE8=0)
    alg_bytes = (node_bytes * counted["node_tests"] + 48 * counted["tri_tests"] + 16 * counted["rays"]) / n + 15.0 * W * rows_local
    k_ms = statistics.mean(kernel_ms) if kernel_ms else ms_per_step
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    traffic, issue = None, None
    traffic_path = os.path.join(ROOT, "profiles", "traffic.json")  # per frame, from the committed ncu --set full capture of the same kernels
    if os.path.exists(traffic_path):
        prof = json.load(open(traffic_path)).get(args.workload, {})
        traffic = prof.get("dram_bytes_per_frame")
        if prof.get("warp_instructions_per_frame") and clocks and clocks.get("sm_mhz") and world == 1:
            # the bound that actually binds (tree + triangles are L2 resident): issue slots = 148 SMs x 4 schedulers x f_SM
            slots = 148 * 4 * clocks["sm_mhz"] * 1e6 * (k_ms * 1e-3)
            issue = {"warp_instructions": prof["warp_instructions_per_frame"], "issue_slots": slots, "frac": prof["warp_instructions_per_frame"] / slots,
                     "source": "smsp__inst_executed.sum of the two walk kernels in " + prof.get("source", "profiles/") + "; slots from the live kernel time and SM clock"}
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "kernel": "closest-hit walk of one frame: k_wf_trace_light (primary level) + k_render_persistent<seeded> (every deeper level, shading included); raygen / level-0 shade / resolve are < 8 % of the frame", "kernel_ms": k_ms, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "node_bytes": node_bytes,
                "issue": issue,
                "per_ray": {"kd_nodes_visited": counted["node_tests"] / max(1, counted["rays"]), "triangle_tests": counted["tri_tests"] / max(1, counted["rays"])}}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            ref = reference_sample(args.workload, 3, 1)
            cpu = {k: ref[k] for k in ("value", "unit", "cores", "kind", "sample")}
        except Exception as exc:  # the baseline is reported, never required for the GPU number
            cpu = {"value": None, "unit": "Mrays/s", "cores": os.cpu_count(), "kind": "reference", "sample": f"failed: {exc}"}

    line = {"metric": "Mrays/s (primary+secondary)", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak" if (frame_parallel or args.partition == "frames") else "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "traversal": "kd", "pipeline": args.pipeline, "partition": ("whole frames per rank (frame k -> rank k % N), no exchange" if frame_parallel else f"row bands of {band_rows}, " + ("pushed into rank 0's frame by peer stores over NVLink + 1-element all-reduce as the barrier" if args.gather == "peer" else "NCCL gather of u8 rgb to rank 0")) if world > 1 else "single GPU",
                       "l2": "256 MiB flush between timed steps", "rays_per_step": job_rays_per_step, "scene_load_s": t_load, "scene_upload_s": t_upload,
                       "scene_bytes": int(flat.n_triangles * 64 + flat.n_nodes * 48 + flat.n_refs * 4)},
            "tiles": tiles, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": 72 + 56, "d2h_bytes_per_step": rows_local * W * 12,
                    "ms_per_step": t_e2e / e2e_steps * 1e3},
            "gpu_launches": (kernel_launches + (1 if peer is not None else 0)) * args.steps}
    print(json.dumps(line), file=_JSON_OUT or sys.stdout, flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--pipeline", default="auto", choices=["auto", "mega", "wavefront", "persistent", "hybrid"])
    ap.add_argument("--partition", default="frames", choices=["frames", "tiles"], help="N > 1 stills: whole frames per rank (weak) or one frame in row bands (strong)")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"], help="N > 1 stills: how the row bands reach rank 0")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
