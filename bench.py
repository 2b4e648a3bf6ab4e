#!/usr/bin/env python
"""bench.py -- the headline benchmark of the render path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c4|c4-diffuse|c1|c3|c3-gi|c5]

A step is one frame of the workload.  Default workload (BASELINE.json configs[3], the one the metric's "1/2/4/8 B200"
is quoted on): the seeded 1 002 528-triangle height field, 3840x2160, max_ray_depth 8, one point light, reflective.

What N > 1 (torchrun, one rank per GPU, scene replicated) measures:
  still workloads   ONE frame cut into interleaved row bands of 8 rows (band b -> rank b % N).  Every rank renders its
                    bands; the quantised bands are pushed into rank 0's frame by peer stores over NVLink and a 1-element
                    all-reduce is the completion barrier -- all inside the timed step.  Total work is fixed: "strong".
  c5 (the orbit)    frame k of the 240-frame camera orbit -> rank k % N: every step each rank renders a DIFFERENT frame
                    and no bytes are exchanged on the device; per-GPU work is fixed: "weak".
Prints ONE JSON line on rank 0.

  value       Mrays/s, whole job, scene resident in HBM, device-timed (CUDA events on the launch stream), max over ranks
  e2e         the same job through the C ABI with HOST buffers, copies inside the timed region: per step the 72-byte
              camera and the parameters go host->device with the launch, and the FLOAT frame (12 B/px, what
              crt::render_image returns) arrives in page-locked host memory -- at N > 1 every rank copies its own bands
              over its own PCIe link to their place in ONE frame in shared host memory (crt_cuda_render_frame), for c5
              every rank delivers its own frames.  e2e.rgb8 is the same with the 3 B/px quantised frame (what write_ppm
              consumes, crt_image_ppm.cpp:16-19).
  roofline    the walk kernels against the measured HBM peak (MEASURED_PEAKS.json) with the ALGORITHMIC bytes
              16*N_node + 48*N_tri + 16 per ray (+ 15 B per pixel), N_node / N_tri counted by the same kernels; and
              roofline.issue: the bound that binds (tree and triangles are L2 resident) -- see DESIGN.md §5
  cpu_baseline the unmodified reference (oracle/_ref) on this box's host cores on a bounded sample (rank 0, N = 1 only)
  sustain     the same frames back to back for >= 2 s: ms per frame and the median SM clock over that time

--impl reference times the reference's own CPU renderer (oracle/_ref/ref_harness -> crt::render_image) on the SAME
workload at the SAME size; it loads nothing of the product (ray counts come from tests/golden/ray_counts.json, which the
GPU suite pins to the device's counters and the CPU suite to the oracle).
"""
from __future__ import annotations

import argparse
import importlib.util
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
BAND_ROWS = 8

WORKLOADS = {
    # name: (description, width, height, max_ray_depth, sample size for the in-line cpu_baseline leg)
    "c4": ("synthetic 1002528-triangle height field, reflective, 3840x2160, max_ray_depth 8, 1 point light", 3840, 2160, 8, (1920, 1080)),
    "c4-diffuse": ("synthetic 1002528-triangle height field, diffuse, 3840x2160, max_ray_depth 8, 1 point light", 3840, 2160, 8, (1920, 1080)),
    "c1": ("scenes/14-01-acceleration-tree/scene1 (== 13-01-optimizations/scene0), 1920x1080, defaults", 1920, 1080, 3, (1920, 1080)),
    "c3": ("scenes/15-01-conclusion/scene1 at 3840x2160, defaults", 3840, 2160, 3, (1920, 1080)),
    "c3-gi": ("scenes/15-01-conclusion/scene2 (GI on, refractive + reflective) at 2160x2160, defaults", 2160, 2160, 3, (360, 360)),
    "c5": ("scenes/11-01-refractive/scene8 1920x1080, one orbit frame per step (240-frame orbit, frame k -> rank k % N)", 1920, 1080, 3, (1920, 1080)),
}
BUNDLED = {"c1": "14-01-acceleration-tree/scene1.crtscene", "c3": "15-01-conclusion/scene1.crtscene", "c3-gi": "15-01-conclusion/scene2.crtscene",
           "c5": "11-01-refractive/scene8.crtscene"}


def scenes_module():
    """chaos-ray-tracing-course-2025_b200/scenes.py on its own (pure numpy): the reference arm must not load the product"""
    if "crt_b200_scenes_standalone" not in sys.modules:
        path = os.path.join(ROOT, "chaos-ray-tracing-course-2025_b200", "scenes.py")
        spec = importlib.util.spec_from_file_location("crt_b200_scenes_standalone", path)
        module = importlib.util.module_from_spec(spec)
        sys.modules["crt_b200_scenes_standalone"] = module
        spec.loader.exec_module(module)
    return sys.modules["crt_b200_scenes_standalone"]


def scene_text(workload: str, size=None) -> tuple[str, str]:
    """(json text, asset root) of the workload, optionally re-sized (CPU sample)."""
    import oracle_tools as ot  # only for where the bundled-scene fixtures lie
    w, h = size or WORKLOADS[workload][1:3]
    if workload.startswith("c4"):
        return scenes_module().height_field_scene(708, w, h, "diffuse" if workload.endswith("diffuse") else "reflective"), ""
    rel = BUNDLED[workload]
    return ot.scene_variant(rel, w, h), os.path.dirname(ot.scene_file(rel))


def golden_counts(workload: str) -> dict:
    return json.load(open(os.path.join(ROOT, "tests", "golden", "ray_counts.json")))[workload]


def orbit_of(text: str, frames: int = 240):
    """the C5 orbit from the scene text alone (camera + vertex bounds): [(position[3], matrix[9])], float32"""
    doc = json.loads(text)
    lo, hi = np.full(3, np.inf, np.float32), np.full(3, -np.inf, np.float32)
    for obj in doc["objects"]:
        v = np.asarray(obj["vertices"], np.float32).reshape(-1, 3)
        if len(v):
            lo, hi = np.minimum(lo, v.min(axis=0)), np.maximum(hi, v.max(axis=0))
    center = (lo.astype(np.float64) + hi.astype(np.float64)) / 2  # the root box of the tree: min / max over all vertices (crt_acceleration_tree.cpp:89-94)
    cam = doc["camera"]
    return scenes_module().orbit_cameras([np.float32(v) for v in cam["position"]], [np.float32(v) for v in cam["matrix"]], center, frames)


def orbit_frame(k: int, frames: int = 240) -> int:
    """the k-th frame a job renders.  The orbit is walked in steps of 97 frames (coprime to 240: every frame comes up once
    in 240) so that a run of 20 steps samples the WHOLE circle whatever N is: a frame costs between 0.15 and 0.65 ms
    depending on where the camera stands (r2l, r2t: the half of the circle a stride of 7 gave a 1-GPU run averages 0.19 ms,
    the frames an 8-rank run saw 0.29 ms -- on the same single GPU), and a scaling figure must not compare two arcs."""
    return (97 * k) % frames


def partition_text(workload: str, world: int, partition: str, gather: str) -> str:
    if world <= 1:
        return "single GPU"
    if partition == "frames":
        return ("orbit: the k-th frame of the job (orbit frame 97k mod 240) -> rank k % N, every rank renders a different frame per step, no exchange" if workload == "c5"
                else "REPLICAS: every rank renders the same whole frame, no exchange (not a scaling measurement of one frame)")
    return f"one frame in interleaved row bands of {BAND_ROWS}, " + ("pushed into rank 0's frame by peer stores over NVLink + 1-element all-reduce as the barrier"
                                                                       if gather == "peer" else "NCCL gather of u8 rgb to rank 0")


def config_for(args, world: int) -> dict:
    """the SAME dict on both arms: what is rendered, at which size, how it is cut over the ranks"""
    desc, W, H, depth, _ = WORKLOADS[args.workload]
    gold = golden_counts(args.workload)
    return {"workload": desc, "width": W, "height": H, "max_ray_depth": depth, "rays_per_frame": gold["rays_per_frame"],
            "partition": partition_text(args.workload, world, args.partition, args.gather), "l2": "256 MiB flush between timed steps"}


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

    def window(self, t0: float, t1: float) -> dict:
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

    def stop(self):
        if self.nvml:
            self.running = False
            self.thread.join(timeout=1.0)
        elif self.proc:
            time.sleep(0.15)
            self.proc.terminate()


def bind_to_gpu_numa_node(local: int) -> dict:
    """run this rank (and first-touch its page-locked buffers) on the host cores next to its GPU"""
    info = {"numa_node": None, "cpus": None}
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local)).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        bus = bus[4:] if len(bus) > 12 else bus  # 00000000:17:00.0 -> 0000:17:00.0
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        info["numa_node"] = node
        if node >= 0:
            cpus = []
            for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
                a, _, b = part.partition("-")
                cpus += list(range(int(a), int(b or a) + 1))
            allowed = sorted(set(cpus) & os.sched_getaffinity(0))
            if allowed:
                os.sched_setaffinity(0, allowed)
                info["cpus"] = len(allowed)
    except Exception as exc:  # a box without NUMA information: nothing to bind
        info["note"] = str(exc)[:80]
    return info


# ------------------------------------------------------------------------------------------------------------- reference
def reference_render(workload: str, frames: int, warmup: int, size=None) -> dict:
    """the reference's own crt::render_image on this box's host cores (oracle/_ref/ref_harness; its own timed region,
    main.cpp:37-43).  c5: each frame is a different orbit frame, each with its own ray count."""
    import oracle_tools as ot
    desc, W, H, depth, _ = WORKLOADS[workload]
    full_size = size is None
    text, root = scene_text(workload, size)
    gold = golden_counts(workload)
    if not ot.have_reference_build():
        return port_render_fallback(workload, frames, warmup, size)
    times, rays = [], []
    with tempfile.TemporaryDirectory() as tmp:
        if workload == "c5":
            orbit = orbit_of(text)
            doc = json.loads(text)
            for k in range(warmup + frames):
                pos, mat = orbit[orbit_frame(k, len(orbit))]
                doc["camera"]["position"] = [float(v) for v in pos]
                doc["camera"]["matrix"] = [float(v) for v in mat]
                path = os.path.join(tmp, f"orbit{k}.crtscene")
                open(path, "w").write(json.dumps(doc))
                out = ot.run_harness(path, asset_root=root or None, max_depth=depth, repeat=2)  # first render of a process: page faults
                assert out.get("loaded"), out
                os.unlink(path)
                if k >= warmup:
                    times.append(out["render_s"][1])
                    rays.append(gold["rays_per_orbit_frame"][orbit_frame(k, len(orbit))] if full_size else None)
        else:
            path = os.path.join(tmp, "scene.crtscene")
            open(path, "w").write(text)
            out = ot.run_harness(path, asset_root=root or None, max_depth=depth, repeat=warmup + frames)
            assert out.get("loaded"), out
            times = out["render_s"][warmup:]
            rays = [gold["rays_per_frame"] if full_size else None] * len(times)
        threads = out["threads"]
    if not full_size:  # a reduced sample: same rays per pixel (same camera, same scene), scaled by the pixel count
        w, h = size
        rays = [gold["rays_per_frame"] * (w * h) / float(W * H)] * len(times)
    sec, n_rays = statistics.mean(times), statistics.mean(rays)
    w, h = size or (W, H)
    return {"value": n_rays / sec / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "reference", "ms_per_step": sec * 1e3, "rays_per_step": n_rays,
            "sample": f"{w}x{h} render of the workload's scene and camera" + ("" if full_size else " (a reduced frame: rays scaled by the pixel count)")
                      + f", {n_rays:.0f} rays/frame, mean of {len(times)} frames, {sec * 1e3:.1f} ms/frame"}


def port_render_fallback(workload: str, frames: int, warmup: int, size) -> dict:
    """oracle/_ref did not travel: the oracle port stands in (kind "port"); it needs the product's flattener"""
    import __graft_entry__ as entry
    import oracle_tools as ot
    crt = entry.load_package()
    depth = WORKLOADS[workload][3]
    text, root = scene_text(workload, size)
    flat = crt.Scene.load_json(text, root).flatten()
    params = crt.default_params(max_ray_depth=depth, traversal=crt.TRAVERSAL_LITERAL)
    times = []
    for _ in range(warmup + frames):
        t0 = time.perf_counter()
        out = ot.port_render(flat, params, primary=False)
        times.append(time.perf_counter() - t0)
    sec, rays = statistics.mean(times[warmup:]), out["stats"]["rays"]
    return {"value": rays / sec / 1e6, "unit": "Mrays/s", "cores": os.cpu_count(), "kind": "port", "ms_per_step": sec * 1e3, "rays_per_step": rays,
            "sample": f"{flat.camera.width}x{flat.camera.height} render by the oracle port ({rays} rays/frame), mean of {frames} frames, {sec * 1e3:.1f} ms/frame"}


def run_reference_arm(args):
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    ref = reference_render(args.workload, args.steps, args.warmup)  # the full-size frame, every step
    line = {"impl": "reference", "metric": "Mrays/s (primary+secondary)", "value": ref["value"], "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ref["ms_per_step"], "higher_is_better": True,
            "scaling": scaling_of(args), "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_for(args, world),
            "cpu_baseline": {"value": ref["value"], "unit": "Mrays/s", "cores": ref["cores"], "kind": ref["kind"], "sample": ref["sample"]},
            "e2e": {"value": ref["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
            "note": "one host, all its threads, whatever N is: the reference has no multi-device path"}
    print(json.dumps(line), flush=True)


def scaling_of(args) -> str:
    return "weak" if args.partition == "frames" else "strong"


# ------------------------------------------------------------------------------------------------------------------ ours
class SharedHostFrame:
    """ONE frame in host memory that every rank of the box has mapped and page-locked: each GPU copies its row bands to
    their place over its own PCIe link (crt_cuda_render_frame).  Each rank touches its own bands first, so their pages
    land on the NUMA node it is bound to."""

    def __init__(self, name: str, nbytes: int, create: bool):
        from multiprocessing import shared_memory
        self.shm = shared_memory.SharedMemory(name=name, create=create, size=nbytes)
        self.array = np.ndarray((nbytes,), dtype=np.uint8, buffer=self.shm.buf)
        self.ptr = self.array.ctypes.data
        self.nbytes, self.creator, self.registered = nbytes, create, False

    def touch(self, byte_ranges):
        for a, b in byte_ranges:
            self.array[a:b] = 0

    def pin(self):
        import torch
        rc = torch.cuda.cudart().cudaHostRegister(self.ptr, self.nbytes, 0)
        if int(rc) != 0:
            raise RuntimeError(f"cudaHostRegister -> {rc}")
        self.registered = True

    def close(self):
        import torch
        if self.registered:
            torch.cuda.cudart().cudaHostUnregister(self.ptr)
        self.array = None
        self.shm.close()
        if self.creator:
            self.shm.unlink()


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    numa = bind_to_gpu_numa_node(local) if not args.no_numa_bind else {"numa_node": None, "cpus": None, "note": "--no-numa-bind"}

    import __graft_entry__ as entry
    crt = entry.load_package()
    from crt_b200 import distributed as D

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

    orbit = orbit_of(text) if args.workload == "c5" else None
    frames_mode = args.partition == "frames"      # whole frames per rank
    tiles_mode = world > 1 and not frames_mode     # one frame in row bands
    pipeline = {"auto": crt.PIPELINE_AUTO, "mega": crt.PIPELINE_MEGAKERNEL, "wavefront": crt.PIPELINE_WAVEFRONT, "persistent": crt.PIPELINE_PERSISTENT,
                "hybrid": crt.PIPELINE_HYBRID}[args.pipeline]
    common = dict(max_ray_depth=depth, traversal=crt.TRAVERSAL_KD, pipeline=pipeline)
    part = dict(rank=rank, world=world, band_rows=BAND_ROWS) if tiles_mode else dict(rank=0, world=1)
    params = crt.default_params(**common, **part)

    def frame_index(step):  # c5: a different orbit frame on every rank, every step
        # ($CRT_ORBIT_AS=rank/world: the frame sequence of that rank of a larger job, on this GPU -- a measurement aid)
        fake = os.environ.get("CRT_ORBIT_AS")
        r_, w_ = (int(v) for v in fake.split("/")) if fake else (rank, world)
        return orbit_frame(step * w_ + r_, len(orbit))

    cameras = {}  # built once: the timed loop should launch kernels, not assemble ctypes structures

    def camera_for(step):
        key = frame_index(step) if orbit is not None else -1
        if key not in cameras:
            cam = crt.Camera.from_buffer_copy(flat.camera)
            if orbit is not None:
                pos, mat = orbit[key]
                cam.position[:] = [float(v) for v in pos]
                cam.rotation[:] = [float(v) for v in mat]
            cameras[key] = cam
        return cameras[key]

    # --- ray / test counts of the timed configuration (counting build of the same kernels, outside the timed region)
    counted = {"rays": 0, "node_tests": 0, "tri_tests": 0, "entered_rays": 0}
    count_steps = list(range(args.warmup, args.warmup + args.steps)) if orbit is not None else [0]
    rays_of_step = {}
    for s in count_steps:
        renderer.set_camera(camera_for(s))
        _, _, st = renderer.render(crt.default_params(collect_stats=1, **common, **part), want_rgb=False, want_rgb8=False)
        rays_of_step[s] = st["rays"]
        for k in counted:
            counted[k] += st[k]
    kernel_launches = st["kernel_launches"]
    n_counted = len(count_steps)

    def all_sum(value):
        t = torch.tensor([float(value)], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t)
        return float(t.item())

    def all_max(value):
        t = torch.tensor([float(value)], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    job_rays_per_step = all_sum(counted["rays"] / n_counted)            # all ranks together: one frame (tiles) or N frames
    job_entered_per_step = all_sum(counted["entered_rays"] / n_counted)

    gather, peer = None, None
    if tiles_mode:
        if args.gather == "peer":
            peer = D.PeerFrame(renderer, W, H, device)  # bands are pushed into rank 0's frame over NVLink (crt_push.cuh)
        else:
            gather = D.BandGather(W, H, BAND_ROWS, 3, device)  # NCCL gather + un-interleave on rank 0
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)  # > 126 MB L2

    def device_step(step):
        renderer.set_camera(camera_for(step))
        renderer.render_device(params)
        if peer is not None:
            peer.push()
        elif gather is not None:
            _, rgb8_ptr, w, rows = renderer.device_framebuffer()
            gather.gather(D.wrap_device_bytes(rgb8_ptr, w * rows * 3, device))

    def timed_steps(n, first_step):
        """n steps, each bracketed by events on the launch stream, L2 flushed before each (outside the bracket); the
        host does NOT wait between steps, so the launches of step s+1 are queued while step s runs: a bracket measures
        the device, not how fast one host thread can issue a dozen launches into an idle GPU (r2c: the 11 launches of a
        C5 frame take the host 0.65 ms that way, the device 0.31 ms).  Returns per-step ms."""
        begin = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
        end = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
        t0 = time.perf_counter()
        for s in range(n):
            flush.fill_(s & 0xFF)
            begin[s].record(stream)
            device_step(first_step + s)
            end[s].record(stream)
        t1 = time.perf_counter()
        torch.cuda.synchronize(device)
        t2 = time.perf_counter()
        renderer.synchronize()  # (the context's own bookkeeping: ray-pool flags of the last frame)
        # how long the host took to ISSUE a step, and a step's share of the wall clock until the device was done: when
        # the two are equal the host is the limit and the device's brackets contain the gaps between its launches
        timed_steps.host = {"issue_ms_per_step": (t1 - t0) / n * 1e3, "wall_ms_per_step": (t2 - t0) / n * 1e3}
        return [b.elapsed_time(e) for b, e in zip(begin, end)]

    sampler = ClockSampler(local) if rank == 0 else None  # (before the barrier: NVML start-up on rank 0 must not delay its first timed step)
    for s in range(args.warmup):
        device_step(s)
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize(device)
    wall0 = time.perf_counter()
    step_ms = timed_steps(args.steps, args.warmup)
    if world > 1:
        dist.barrier()
    wall1 = time.perf_counter()
    total_ms = all_max(sum(step_ms))
    clocks = sampler.window(wall0, wall1) if sampler else None
    per_rank = [None] * world  # every rank's own view of its steps (the line reports the slowest rank)
    mine = {"rank": rank, **timed_steps.host, "ms_mean": statistics.mean(step_ms), "ms_min": min(step_ms), "ms_median": statistics.median(step_ms), "ms_max": max(step_ms)}
    if world > 1:
        dist.all_gather_object(per_rank, mine)
    else:
        per_rank = [mine]

    # --- sustained: the same steps back to back for >= args.sustain seconds (does the clock of the 60 ms region hold?)
    sustain = None
    if args.sustain > 0 and world == 1:
        n_sus = max(args.steps, int(args.sustain * 1e3 / max(1e-3, sum(step_ms) / len(step_ms) + 0.15)))
        s0 = time.perf_counter()
        sus_ms = timed_steps(n_sus, args.warmup)
        s1 = time.perf_counter()
        sus_clocks = sampler.window(s0, s1) if sampler else None
        sustain = {"frames": n_sus, "seconds": s1 - s0, "ms_per_step": statistics.mean(sus_ms), "ms_per_step_last_quarter": statistics.mean(sus_ms[-max(1, n_sus // 4):]),
                   "clocks": sus_clocks}

    # --- e2e: the C-ABI call with HOST buffers (page-locked), copies inside the timed region
    def e2e_run(kind):
        """kind 'rgb' (12 B/px floats) or 'rgb8' (3 B/px).  Returns (seconds for args.steps steps (max over ranks), d2h bytes per step of the whole job)"""
        bpp = 12 if kind == "rgb" else 3
        frame_bytes = W * H * bpp
        shared = None
        if tiles_mode:
            name = [f"crt_bench_{os.getpid()}_{kind}"] if rank == 0 else [None]
            dist.broadcast_object_list(name, src=0)
            if rank == 0:
                shared = SharedHostFrame(name[0], frame_bytes, create=True)
            dist.barrier()
            if rank != 0:
                shared = SharedHostFrame(name[0], frame_bytes, create=False)
            rows = crt.Renderer.owned_rows(H, rank, world, BAND_ROWS)
            starts = rows[::BAND_ROWS]
            shared.touch([(int(r) * W * bpp, int(min(H, r + BAND_ROWS)) * W * bpp) for r in starts])
            shared.pin()
            host_ptr, d2h = shared.ptr, frame_bytes
        else:
            host = torch.empty(frame_bytes, dtype=torch.uint8).pin_memory()
            host_ptr, d2h = host.data_ptr(), frame_bytes * (world if frames_mode else 1)

        def step(s):
            renderer.set_camera(camera_for(s))
            if kind == "rgb":
                renderer.render_frame(params, rgb_frame=host_ptr)
            else:
                renderer.render_frame(params, rgb8_frame=host_ptr)

        for s in range(min(3, args.warmup)):
            step(s)
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for s in range(args.steps):
            step(args.warmup + s)
        if tiles_mode:
            dist.barrier()  # the frame is complete when every rank's bands have landed
        seconds = all_max(time.perf_counter() - t0)
        if shared is not None:
            dist.barrier()
            shared.close()
        return seconds, d2h

    t_e2e, d2h_rgb = e2e_run("rgb")
    t_e2e8, d2h_rgb8 = e2e_run("rgb8")

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    if sampler:
        sampler.stop()

    ms_per_step = total_ms / args.steps
    value = job_rays_per_step / (ms_per_step * 1e-3) / 1e6
    e2e_value = job_rays_per_step * args.steps / t_e2e / 1e6
    e2e8_value = job_rays_per_step * args.steps / t_e2e8 / 1e6

    # --- roofline of the walk kernels of one frame on this rank
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    else:
        peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
    rows_local = len(crt.Renderer.owned_rows(H, params.rank, params.world, params.band_rows))
    node_bytes = crt.lib.crt_cuda_walk_node_bytes()  # 16: crt_kd_node as it is (8 with -DCRT_NODE8=1)
    alg_bytes = (node_bytes * counted["node_tests"] + 48 * counted["tri_tests"] + 16 * counted["rays"]) / n_counted + 15.0 * W * rows_local
    # the render kernels of one frame on this rank: the step's own bracket (at N > 1 it also holds the band push and the barrier)
    k_ms = statistics.mean(step_ms)
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    prof = {}
    traffic_path = os.path.join(ROOT, "profiles", "traffic.json")  # per frame, from the committed ncu --set full capture of the same kernels
    if os.path.exists(traffic_path):
        prof = json.load(open(traffic_path)).get(args.workload, {})
    issue = None
    if clocks and clocks.get("sm_mhz") and world == 1:
        slots = 148 * 4 * clocks["sm_mhz"] * 1e6 * (k_ms * 1e-3)  # warp-instruction issue slots of the live kernel time
        issue = {"issue_slots": slots}
        if prof.get("inst_alg"):
            # inst_alg = the kernel's own NON-DIVERGENT warp instructions per node visit / triangle test / ray,
            # measured with ncu on a fully coherent ray batch (profiles/, tools/inst_alg.py); 32 rays share one instruction
            ia = prof["inst_alg"]
            walk_visits = counted["node_tests"] - counted["rays"]  # (one "node test" per ray is the root-box test of the ray set-up)
            alg_inst = (ia["per_node_visit"] * walk_visits + ia["per_triangle_test"] * counted["tri_tests"]
                        + ia.get("per_ray", 0.0) * counted["entered_rays"]) / n_counted / 32.0
            issue.update({"bound": "issue", "algorithmic_warp_instructions": alg_inst, "frac": alg_inst / slots, "inst_alg": ia,
                          "definition": "(inst_alg,node * N_node + inst_alg,tri * N_tri + inst_alg,ray * N_ray) / 32 lanes / (148 SM x 4 schedulers x f_SM x kernel time)"})
        if prof.get("warp_instructions_per_frame"):
            issue["issue_slot_utilisation"] = prof["warp_instructions_per_frame"] / slots  # executed (divergent) warp instructions, not a bound
            issue["executed_warp_instructions"] = prof["warp_instructions_per_frame"]
            issue["source"] = prof.get("source", "profiles/")
    n_rays = counted["rays"] / n_counted
    n_entered = max(1.0, counted["entered_rays"] / n_counted)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": prof.get("dram_bytes_per_frame"),
                "kernel": prof.get("kernel", "the closest-hit walk kernels of one frame (k_wf_trace_lean per wavefront level, k_render_persistent for the deeper levels / GI)"),
                "kernel_ms": k_ms, "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes, "node_bytes": node_bytes, "issue": issue,
                "per_ray": {"kd_nodes_visited": counted["node_tests"] / max(1.0, counted["rays"]), "triangle_tests": counted["tri_tests"] / max(1.0, counted["rays"])},
                "per_traversing_ray": {"kd_nodes_visited": counted["node_tests"] / n_counted / n_entered, "triangle_tests": counted["tri_tests"] / n_counted / n_entered,
                                       "note": "rays that enter the root box; the others end in ray setup (one slab test)"}}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            sample = WORKLOADS[args.workload][4]
            ref = reference_render(args.workload, 3, 1, size=None if tuple(sample) == (W, H) else sample)
            cpu = {k: ref[k] for k in ("value", "unit", "cores", "kind", "sample")}
        except Exception as exc:  # the baseline is reported, never required for the GPU number
            cpu = {"value": None, "unit": "Mrays/s", "cores": os.cpu_count(), "kind": "reference", "sample": f"failed: {exc}"}

    config = config_for(args, world)
    gold = golden_counts(args.workload)
    mean_gold = (statistics.mean(gold["rays_per_orbit_frame"][frame_index(s)] for s in count_steps) if orbit is not None and "rays_per_orbit_frame" in gold
                 else gold["rays_per_frame"])
    line = {"metric": "Mrays/s (primary+secondary)", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": scaling_of(args), "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config,
            "details": {"traversal": "kd", "pipeline": args.pipeline, "rays_per_step": job_rays_per_step, "traversing_rays_per_step": job_entered_per_step,
                        "Mrays_per_s_traversing": job_entered_per_step / (ms_per_step * 1e-3) / 1e6,
                        "rays_per_frame_counted_on_this_rank": n_rays, "rays_per_frame_golden": mean_gold,
                        "scene_load_s": t_load, "scene_upload_s": t_upload, "scene_bytes": int(flat.n_triangles * 64 + flat.n_fast_nodes * 16 + flat.n_fast_refs * 4),
                        "numa": numa, "frames_per_step": world if frames_mode and world > 1 else 1, "per_rank_step_ms": per_rank},
            "clocks": clocks, "sustain": sustain, "roofline": roofline, "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": (72 + 56) * world, "d2h_bytes_per_step": d2h_rgb,
                    "ms_per_step": t_e2e / args.steps * 1e3,
                    "what": "float frame (12 B/px, what crt::render_image returns) in page-locked host memory" + (": every rank's bands copied to their place in one shared frame" if tiles_mode else ""),
                    "rgb8": {"value": e2e8_value, "unit": "Mrays/s", "d2h_bytes_per_step": d2h_rgb8, "ms_per_step": t_e2e8 / args.steps * 1e3,
                             "what": "quantised frame (3 B/px, what write_ppm consumes)"}},
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
    ap.add_argument("--no-numa-bind", action="store_true")
    ap.add_argument("--sustain", type=float, default=2.0, help="seconds of back-to-back frames after the timed steps (N = 1); 0 = skip")
    ap.add_argument("--pipeline", default="auto", choices=["auto", "mega", "wavefront", "persistent", "hybrid"])
    ap.add_argument("--partition", default=None, choices=["frames", "tiles"],
                    help="N > 1: one frame in row bands (tiles: the default for stills) or whole frames per rank (frames: the default for c5)")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"], help="tiles: how the row bands reach rank 0")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.partition is None:
        args.partition = "frames" if args.workload == "c5" else "tiles"
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
