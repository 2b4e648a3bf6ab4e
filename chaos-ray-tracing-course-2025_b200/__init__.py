"""ctypes binding of the C ABI in include/crt_cuda.h (lib/libcrt_b200.so) plus a Python mirror of the
reference's operator interface (crt::render_image / RendererSettings, src/core/crt_renderer.h:18-27).

Nothing here computes pixels: every result comes from the CUDA kernels behind the C ABI.  If the shared library
is missing the import fails loudly; if there is no CUDA device, creating a Renderer raises -- there is no CPU path.
The directory name contains hyphens, so load the package with ``__graft_entry__.load_package()`` (alias ``crt_b200``).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# CRT_B200_LIB: a differently tuned build of the SAME library (kernel parameter sweeps); never a different backend
LIB_PATH = os.environ.get("CRT_B200_LIB") or os.path.join(_HERE, "lib", "libcrt_b200.so")

CRT_OK = 0
ERR_INVALID, ERR_CUDA, ERR_IO, ERR_PARSE, ERR_NO_DEVICE = -1, -2, -3, -4, -5
TRAVERSAL_LITERAL, TRAVERSAL_KD, TRAVERSAL_BRUTE = 0, 1, 2
FLATTEN_REFERENCE_TREE = 1
PIPELINE_AUTO, PIPELINE_MEGAKERNEL, PIPELINE_WAVEFRONT, PIPELINE_PERSISTENT, PIPELINE_HYBRID = 0, 1, 2, 3, 4
REF_CULL_BIT = 0x80000000
PPM_NOALLOC_MAX_PIXELS = 8388608  # CRT_PPM_NOALLOC_MAX_PIXELS (include/crt_cuda.h)

# every symbol include/crt_cuda.h declares (tests check that the library exports each of them)
EXPORTED_SYMBOLS = [
    "crt_scene_load_file", "crt_scene_load_json", "crt_scene_flatten", "crt_scene_flatten_ex", "crt_scene_set_camera", "crt_scene_set_resolution",
    "crt_scene_free", "crt_scene_last_error", "crt_render_params_default", "crt_cuda_device_count", "crt_cuda_create",
    "crt_cuda_destroy", "crt_cuda_last_error", "crt_cuda_set_stream", "crt_cuda_upload_scene", "crt_cuda_pack_walk_tree", "crt_cuda_walk_node_bytes", "crt_cuda_set_camera", "crt_cuda_update_shading", "crt_cuda_group_update_shading",
    "crt_cuda_render", "crt_cuda_render_frame", "crt_cuda_render_device", "crt_cuda_device_framebuffer", "crt_cuda_synchronize",
    "crt_cuda_primary_hits", "crt_cuda_primary_rays", "crt_cuda_encode_ppm", "crt_cuda_encode_ppm_device", "crt_cuda_encode_ppm_noalloc",
    "crt_cuda_frame_create", "crt_cuda_frame_open", "crt_cuda_frame_attach", "crt_cuda_frame_pointer", "crt_cuda_push_bands",
    "crt_cuda_group_create", "crt_cuda_group_destroy", "crt_cuda_group_last_error", "crt_cuda_group_size", "crt_cuda_group_upload_scene",
    "crt_cuda_group_set_camera", "crt_cuda_group_render", "crt_cuda_group_frame", "crt_cuda_group_encode_ppm",
]


class Camera(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("rotation", C.c_float * 9), ("tan_half_fov", C.c_float), ("aspect", C.c_float),
                ("width", C.c_int32), ("height", C.c_int32)]


class FlatScene(C.Structure):
    _fields_ = [
        ("camera", Camera), ("background", C.c_float * 3),
        ("gi_on", C.c_int32), ("reflections_on", C.c_int32), ("refractions_on", C.c_int32), ("bucket_size", C.c_int32),
        ("root_min", C.c_float * 3), ("root_max", C.c_float * 3),
        ("n_triangles", C.c_int32), ("n_vertices", C.c_int32), ("n_nodes", C.c_int32), ("n_refs", C.c_int32),
        ("n_lights", C.c_int32), ("n_materials", C.c_int32), ("n_textures", C.c_int32), ("n_texels", C.c_int32),
        ("max_depth", C.c_int32), ("fast_max_depth", C.c_int32), ("n_fast_nodes", C.c_int32), ("n_fast_refs", C.c_int32),
        ("tri_verts", C.c_void_p), ("tri_meta", C.c_void_p), ("vert_normal", C.c_void_p), ("vert_uv", C.c_void_p),
        ("box_nodes", C.c_void_p), ("kd_nodes", C.c_void_p), ("refs", C.c_void_p),
        ("fast_nodes", C.c_void_p), ("fast_refs", C.c_void_p), ("lights", C.c_void_p),
        ("materials", C.c_void_p), ("textures", C.c_void_p), ("texels", C.c_void_p),
    ]

    def _view(self, ptr, count, dtype):
        if not ptr or count == 0:
            return np.zeros(0, dtype=dtype)
        nbytes = count * np.dtype(dtype).itemsize
        buf = (C.c_char * nbytes).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype, count=count)

    def arrays(self):
        """numpy views (no copy) of every flattened array; valid while the owning Scene lives."""
        return {
            "tri_verts": self._view(self.tri_verts, self.n_triangles * 12, np.float32).reshape(-1, 3, 4),
            "tri_meta": self._view(self.tri_meta, self.n_triangles * 4, np.int32).reshape(-1, 4),
            "vert_normal": self._view(self.vert_normal, self.n_vertices * 4, np.float32).reshape(-1, 4),
            "vert_uv": self._view(self.vert_uv, self.n_vertices * 2, np.float32).reshape(-1, 2),
            "box_nodes": self._view(self.box_nodes, self.n_nodes * 8, np.int32).reshape(-1, 8),
            "kd_nodes": self._view(self.kd_nodes, self.n_nodes * 4, np.int32).reshape(-1, 4),
            "refs": self._view(self.refs, self.n_refs, np.uint32),
            "fast_nodes": self._view(self.fast_nodes, self.n_fast_nodes * 4, np.int32).reshape(-1, 4),
            "fast_refs": self._view(self.fast_refs, self.n_fast_refs, np.uint32),
            "lights": self._view(self.lights, self.n_lights * 4, np.float32).reshape(-1, 4),
            "materials": self._view(self.materials, self.n_materials * 4, np.int32).reshape(-1, 4),
            "textures": self._view(self.textures, self.n_textures * 12, np.int32).reshape(-1, 12),
            "texels": self._view(self.texels, self.n_texels * 4, np.float32).reshape(-1, 4),
        }


class RenderParams(C.Structure):
    _fields_ = [("max_ray_depth", C.c_uint32), ("diffuse_reflection_ray_count", C.c_uint32),
                ("shadow_bias", C.c_float), ("reflection_bias", C.c_float), ("diffuse_reflection_bias", C.c_float),
                ("refraction_bias", C.c_float),
                ("traversal", C.c_int32), ("pipeline", C.c_int32), ("shadows", C.c_int32), ("collect_stats", C.c_int32),
                ("rank", C.c_int32), ("world", C.c_int32), ("band_rows", C.c_int32), ("pad", C.c_int32)]


class RenderStats(C.Structure):
    _fields_ = [("gpu_ms", C.c_float), ("kernel_launches", C.c_int32), ("rows_rendered", C.c_int32), ("pad", C.c_int32),
                ("rays", C.c_uint64), ("rays_by_depth", C.c_uint64 * 16), ("shadow_rays", C.c_uint64),
                ("node_tests", C.c_uint64), ("tri_tests", C.c_uint64), ("entered_rays", C.c_uint64)]

    def as_dict(self):
        return {"gpu_ms": self.gpu_ms, "kernel_launches": self.kernel_launches, "rows_rendered": self.rows_rendered,
                "rays": self.rays, "rays_by_depth": list(self.rays_by_depth), "shadow_rays": self.shadow_rays,
                "node_tests": self.node_tests, "tri_tests": self.tri_tests, "entered_rays": self.entered_rays}


def _load_library():
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(the product path is the CUDA library; there is no Python or CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    P = C.POINTER
    vp = C.c_void_p
    sig = {
        "crt_scene_load_file": (C.c_int, [C.c_char_p, P(vp)]),
        "crt_scene_load_json": (C.c_int, [C.c_char_p, C.c_size_t, C.c_char_p, P(vp)]),
        "crt_scene_flatten": (C.c_int, [vp, P(FlatScene)]),
        "crt_scene_flatten_ex": (C.c_int, [vp, C.c_int, P(FlatScene)]),
        "crt_cuda_pack_walk_tree": (C.c_int, [vp, C.c_int32, vp]),
        "crt_cuda_walk_node_bytes": (C.c_int, []),
        "crt_scene_set_camera": (C.c_int, [vp, P(C.c_float), P(C.c_float), C.c_float]),
        "crt_scene_set_resolution": (C.c_int, [vp, C.c_int, C.c_int]),
        "crt_scene_free": (None, [vp]),
        "crt_scene_last_error": (C.c_char_p, []),
        "crt_render_params_default": (None, [P(RenderParams)]),
        "crt_cuda_device_count": (C.c_int, []),
        "crt_cuda_create": (C.c_int, [C.c_int, P(vp)]),
        "crt_cuda_destroy": (None, [vp]),
        "crt_cuda_last_error": (C.c_char_p, [vp]),
        "crt_cuda_set_stream": (C.c_int, [vp, vp]),
        "crt_cuda_upload_scene": (C.c_int, [vp, P(FlatScene)]),
        "crt_cuda_set_camera": (C.c_int, [vp, P(Camera)]),
        "crt_cuda_update_shading": (C.c_int, [vp, P(FlatScene)]),
        "crt_cuda_group_update_shading": (C.c_int, [vp, P(FlatScene)]),
        "crt_cuda_render": (C.c_int, [vp, P(RenderParams), vp, vp, P(RenderStats)]),
        "crt_cuda_render_frame": (C.c_int, [vp, P(RenderParams), vp, vp, P(RenderStats)]),
        "crt_cuda_render_device": (C.c_int, [vp, P(RenderParams), P(RenderStats)]),
        "crt_cuda_device_framebuffer": (C.c_int, [vp, P(vp), P(vp), P(C.c_int), P(C.c_int)]),
        "crt_cuda_synchronize": (C.c_int, [vp, P(RenderStats)]),
        "crt_cuda_primary_hits": (C.c_int, [vp, C.c_int, vp, vp, vp, vp, P(RenderStats)]),
        "crt_cuda_primary_rays": (C.c_int, [vp, vp]),
        "crt_cuda_encode_ppm": (C.c_int, [vp, vp, C.c_size_t, P(C.c_size_t)]),
        "crt_cuda_encode_ppm_device": (C.c_int, [vp, vp, C.c_int, C.c_int, vp, C.c_size_t, P(C.c_size_t)]),
        "crt_cuda_encode_ppm_noalloc": (C.c_int, [vp, vp, C.c_int, C.c_int, vp, C.c_size_t, P(C.c_size_t)]),
        "crt_cuda_frame_create": (C.c_int, [vp, C.c_int, C.c_int, vp]),
        "crt_cuda_frame_open": (C.c_int, [vp, vp, C.c_int, C.c_int]),
        "crt_cuda_frame_attach": (C.c_int, [vp, vp, C.c_int, C.c_int]),
        "crt_cuda_frame_pointer": (C.c_int, [vp, P(vp), P(C.c_int), P(C.c_int)]),
        "crt_cuda_push_bands": (C.c_int, [vp]),
        "crt_cuda_group_create": (C.c_int, [P(C.c_int), C.c_int, P(vp)]),
        "crt_cuda_group_destroy": (None, [vp]),
        "crt_cuda_group_last_error": (C.c_char_p, [vp]),
        "crt_cuda_group_size": (C.c_int, [vp]),
        "crt_cuda_group_upload_scene": (C.c_int, [vp, P(FlatScene)]),
        "crt_cuda_group_set_camera": (C.c_int, [vp, P(Camera)]),
        "crt_cuda_group_render": (C.c_int, [vp, P(RenderParams), vp, vp, P(RenderStats)]),
        "crt_cuda_group_frame": (C.c_int, [vp, P(vp), P(C.c_int), P(C.c_int)]),
        "crt_cuda_group_encode_ppm": (C.c_int, [vp, vp, C.c_size_t, P(C.c_size_t)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load_library()


class CrtError(RuntimeError):
    pass


def default_params(**overrides) -> RenderParams:
    p = RenderParams()
    lib.crt_render_params_default(C.byref(p))
    for k, v in overrides.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


class Scene:
    """Host scene handle: parse + tree build + flatten (crt::json::read_scene_from_istream, src/core/crt_json.h:11)."""

    def __init__(self, handle):
        self._h = handle
        self._flat = None

    @classmethod
    def load_file(cls, path: str) -> "Scene":
        h = C.c_void_p()
        rc = lib.crt_scene_load_file(os.fsencode(path), C.byref(h))
        if rc != CRT_OK:
            raise CrtError(f"crt_scene_load_file({path}) -> {rc}: {lib.crt_scene_last_error().decode()}")
        return cls(h)

    @classmethod
    def load_json(cls, text: str | bytes, asset_root: str = "") -> "Scene":
        data = text.encode() if isinstance(text, str) else text
        h = C.c_void_p()
        rc = lib.crt_scene_load_json(data, len(data), os.fsencode(asset_root), C.byref(h))
        if rc != CRT_OK:
            raise CrtError(f"crt_scene_load_json -> {rc}: {lib.crt_scene_last_error().decode()}")
        return cls(h)

    def flatten(self, reference_tree: bool = True) -> FlatScene:
        """reference_tree=False (crt_scene_flatten_ex without CRT_FLATTEN_REFERENCE_TREE): neither builds nor uploads
        the reference's acceleration tree; the scene then renders with TRAVERSAL_KD only"""
        flat = FlatScene()
        rc = lib.crt_scene_flatten_ex(self._h, FLATTEN_REFERENCE_TREE if reference_tree else 0, C.byref(flat))
        if rc != CRT_OK:
            raise CrtError(f"crt_scene_flatten_ex -> {rc}")
        flat._owner = self  # the arrays live in the scene handle: keep it alive as long as the view
        self._flat = flat
        return flat

    def set_resolution(self, width: int, height: int):
        if lib.crt_scene_set_resolution(self._h, width, height) != CRT_OK:
            raise CrtError("crt_scene_set_resolution failed")

    def set_camera(self, position, rotation, fov_degrees: float = 0.0):
        pos = (C.c_float * 3)(*[float(v) for v in position])
        rot = (C.c_float * 9)(*[float(v) for v in np.asarray(rotation, dtype=np.float64).reshape(-1)])
        if lib.crt_scene_set_camera(self._h, pos, rot, float(fov_degrees)) != CRT_OK:
            raise CrtError("crt_scene_set_camera failed")

    def close(self):
        if self._h:
            lib.crt_scene_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Renderer:
    """Device context (crt_cuda_ctx).  Raises when no CUDA device is available."""

    def __init__(self, device: int = 0):
        self._ctx = C.c_void_p()
        rc = lib.crt_cuda_create(device, C.byref(self._ctx))
        if rc != CRT_OK:
            self._ctx = None
            raise CrtError(f"crt_cuda_create({device}) -> {rc} (no CUDA device? this renderer has no CPU fallback)")
        self.width = self.height = 0

    def _check(self, rc, what):
        if rc != CRT_OK:
            raise CrtError(f"{what} -> {rc}: {lib.crt_cuda_last_error(self._ctx).decode()}")

    def set_stream(self, stream_ptr: int | None):
        """None = the context's own stream; 0 (torch's default stream) is passed as cudaStreamLegacy (handle 1)."""
        if stream_ptr is None:
            handle = 0
        else:
            handle = stream_ptr if stream_ptr != 0 else 1
        self._check(lib.crt_cuda_set_stream(self._ctx, C.c_void_p(handle)), "crt_cuda_set_stream")

    def upload(self, scene: Scene | FlatScene):
        flat = scene.flatten() if isinstance(scene, Scene) else scene
        self._check(lib.crt_cuda_upload_scene(self._ctx, C.byref(flat)), "crt_cuda_upload_scene")
        self.width, self.height = flat.camera.width, flat.camera.height
        return flat

    def update_shading(self, flat: FlatScene):
        """lights / materials / texture records / background / flags of `flat` replace the uploaded ones (same counts)"""
        self._check(lib.crt_cuda_update_shading(self._ctx, C.byref(flat)), "crt_cuda_update_shading")

    def set_camera(self, camera: Camera):
        self._check(lib.crt_cuda_set_camera(self._ctx, C.byref(camera)), "crt_cuda_set_camera")
        self.width, self.height = camera.width, camera.height

    @staticmethod
    def owned_rows(height: int, rank: int, world: int, band_rows: int):
        """image rows (increasing y) that `rank` owns under band ownership"""
        if world <= 1:
            return np.arange(height)
        bands = (height + band_rows - 1) // band_rows
        rows = [np.arange(b * band_rows, min(height, (b + 1) * band_rows)) for b in range(rank, bands, world)]
        return np.concatenate(rows) if rows else np.zeros(0, dtype=np.int64)

    def render(self, params: RenderParams | None = None, want_rgb=True, want_rgb8=True, rgb_out=None, rgb8_out=None):
        """crt_cuda_render with host buffers.  Returns (rgb float32 [rows,W,3] | None, rgb8 uint8 | None, stats dict)."""
        params = params or default_params()
        rows = len(self.owned_rows(self.height, params.rank, params.world, params.band_rows))
        rgb = rgb_out if rgb_out is not None else (np.empty((rows, self.width, 3), np.float32) if want_rgb else None)
        rgb8 = rgb8_out if rgb8_out is not None else (np.empty((rows, self.width, 3), np.uint8) if want_rgb8 else None)
        stats = RenderStats()
        rc = lib.crt_cuda_render(self._ctx, C.byref(params), _ptr(rgb), _ptr(rgb8), C.byref(stats))
        self._check(rc, "crt_cuda_render")
        return rgb, rgb8, stats.as_dict()

    def render_frame(self, params: RenderParams, rgb_frame: int | None = None, rgb8_frame: int | None = None):
        """crt_cuda_render_frame: this rank's row bands go to their place in a FULL frame in host memory (raw pointers,
        e.g. of a page-locked shared-memory segment every rank has mapped).  Returns the stats dict."""
        stats = RenderStats()
        rc = lib.crt_cuda_render_frame(self._ctx, C.byref(params), C.c_void_p(rgb_frame), C.c_void_p(rgb8_frame), C.byref(stats))
        self._check(rc, "crt_cuda_render_frame")
        return stats.as_dict()

    def render_device(self, params: RenderParams | None = None):
        params = params or default_params()
        self._check(lib.crt_cuda_render_device(self._ctx, C.byref(params), None), "crt_cuda_render_device")

    def synchronize(self):
        stats = RenderStats()
        self._check(lib.crt_cuda_synchronize(self._ctx, C.byref(stats)), "crt_cuda_synchronize")
        return stats.as_dict()

    def device_framebuffer(self):
        rgb, rgb8, w, r = C.c_void_p(), C.c_void_p(), C.c_int(), C.c_int()
        self._check(lib.crt_cuda_device_framebuffer(self._ctx, C.byref(rgb), C.byref(rgb8), C.byref(w), C.byref(r)), "crt_cuda_device_framebuffer")
        return rgb.value, rgb8.value, w.value, r.value

    def primary_hits(self, traversal: int = TRAVERSAL_KD, counters: bool = True):
        n = (self.height, self.width)
        tri = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        nn = np.empty(n, np.int32) if counters else None
        nt = np.empty(n, np.int32) if counters else None
        stats = RenderStats()
        rc = lib.crt_cuda_primary_hits(self._ctx, traversal, _ptr(tri), _ptr(t), _ptr(nn), _ptr(nt), C.byref(stats))
        self._check(rc, "crt_cuda_primary_hits")
        return tri, t, nn, nt, stats.as_dict()

    def primary_rays(self):
        rays = np.empty((self.height, self.width, 6), np.float32)
        self._check(lib.crt_cuda_primary_rays(self._ctx, _ptr(rays)), "crt_cuda_primary_rays")
        return rays

    def encode_ppm(self) -> bytes:
        length = C.c_size_t()
        self._check(lib.crt_cuda_encode_ppm(self._ctx, None, 0, C.byref(length)), "crt_cuda_encode_ppm")
        buf = C.create_string_buffer(length.value)
        self._check(lib.crt_cuda_encode_ppm(self._ctx, buf, length.value, C.byref(length)), "crt_cuda_encode_ppm")
        return buf.raw[: length.value]

    # ---- multi-GPU stills: the frame on rank 0 that every rank pushes its row bands into (include/crt_cuda.h)
    def frame_create(self, width: int, height: int) -> bytes:
        handle = C.create_string_buffer(64)
        self._check(lib.crt_cuda_frame_create(self._ctx, width, height, handle), "crt_cuda_frame_create")
        return handle.raw

    def frame_open(self, handle: bytes, width: int, height: int):
        self._check(lib.crt_cuda_frame_open(self._ctx, C.create_string_buffer(handle, 64), width, height), "crt_cuda_frame_open")

    def frame_attach(self, frame_ptr: int, width: int, height: int):
        self._check(lib.crt_cuda_frame_attach(self._ctx, C.c_void_p(frame_ptr), width, height), "crt_cuda_frame_attach")

    def frame_pointer(self):
        ptr, w, h = C.c_void_p(), C.c_int(), C.c_int()
        self._check(lib.crt_cuda_frame_pointer(self._ctx, C.byref(ptr), C.byref(w), C.byref(h)), "crt_cuda_frame_pointer")
        return ptr.value, w.value, h.value

    def push_bands(self):
        self._check(lib.crt_cuda_push_bands(self._ctx), "crt_cuda_push_bands")

    def encode_ppm_device(self, rgb8_ptr: int, width: int, height: int) -> bytes:
        """P3 text of any quantised frame in device memory (e.g. the gathered frame on rank 0)"""
        length = C.c_size_t()
        self._check(lib.crt_cuda_encode_ppm_device(self._ctx, C.c_void_p(rgb8_ptr), width, height, None, 0, C.byref(length)), "crt_cuda_encode_ppm_device")
        buf = C.create_string_buffer(length.value)
        self._check(lib.crt_cuda_encode_ppm_device(self._ctx, C.c_void_p(rgb8_ptr), width, height, buf, length.value, C.byref(length)), "crt_cuda_encode_ppm_device")
        return buf.raw[: length.value]

    def encode_ppm_noalloc(self, rgb8_ptr: int, width: int, height: int, text_ptr: int, capacity: int) -> int:
        """P3 text of a quantised device frame written into the CALLER's device buffer; allocates nothing (static device
        scratch, at most PPM_NOALLOC_MAX_PIXELS pixels; calls that overlap on the GPU race).  Returns the text length;
        text_ptr = 0 only queries it."""
        length = C.c_size_t()
        self._check(lib.crt_cuda_encode_ppm_noalloc(self._ctx, C.c_void_p(rgb8_ptr), width, height, C.c_void_p(text_ptr or None), capacity, C.byref(length)),
                    "crt_cuda_encode_ppm_noalloc")
        return length.value

    def close(self):
        if self._ctx:
            lib.crt_cuda_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class RendererGroup:
    """One host thread, several GPUs (crt_cuda_group_*): a context per device, scene replicated, the frame cut into
    interleaved row bands that every device pushes into the frame on devices[0]."""

    def __init__(self, devices):
        self._g = C.c_void_p()
        arr = (C.c_int * len(devices))(*devices)
        rc = lib.crt_cuda_group_create(arr, len(devices), C.byref(self._g))
        if rc != CRT_OK:
            raise CrtError(f"crt_cuda_group_create failed ({rc}); the render path needs CUDA devices with peer access, there is no CPU fallback")
        self.width = self.height = 0

    def _check(self, rc, what):
        if rc != CRT_OK:
            raise CrtError(f"{what} failed ({rc}): {lib.crt_cuda_group_last_error(self._g).decode()}")

    def upload(self, scene):
        flat = scene.flatten() if isinstance(scene, Scene) else scene
        self._check(lib.crt_cuda_group_upload_scene(self._g, C.byref(flat)), "crt_cuda_group_upload_scene")
        self.width, self.height = flat.camera.width, flat.camera.height
        return flat

    def render(self, params: RenderParams | None = None, want_rgb=True, want_rgb8=True):
        params = params or default_params()
        rgb = np.empty((self.height, self.width, 3), np.float32) if want_rgb else None
        rgb8 = np.empty((self.height, self.width, 3), np.uint8) if want_rgb8 else None
        stats = RenderStats()
        self._check(lib.crt_cuda_group_render(self._g, C.byref(params), _ptr(rgb), _ptr(rgb8), C.byref(stats)), "crt_cuda_group_render")
        return rgb, rgb8, stats.as_dict()

    def encode_ppm(self) -> bytes:
        length = C.c_size_t()
        self._check(lib.crt_cuda_group_encode_ppm(self._g, None, 0, C.byref(length)), "crt_cuda_group_encode_ppm")
        buf = C.create_string_buffer(length.value)
        self._check(lib.crt_cuda_group_encode_ppm(self._g, buf, length.value, C.byref(length)), "crt_cuda_group_encode_ppm")
        return buf.raw[: length.value]

    def close(self):
        if self._g:
            lib.crt_cuda_group_destroy(self._g)
            self._g = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return C.c_void_p(a.ctypes.data)
    return C.c_void_p(int(a))  # raw address (e.g. a pinned torch tensor's data_ptr())


# ---------------------------------------------------------------------------------------------------------------
# Python mirror of the reference operator interface (src/core/crt_renderer.h:18-27)
# ---------------------------------------------------------------------------------------------------------------
DEFAULT_SCENE_BUCKET_SIZE = 24
DEFAULT_MAX_RAY_DEPTH = 3
DEFAULT_DIFFUSE_REFLECTION_RAY_COUNT = 4
DEFAULT_SHADOW_BIAS = DEFAULT_REFLECTION_BIAS = DEFAULT_DIFFUSE_REFLECTION_BIAS = DEFAULT_REFRACTION_BIAS = 1e-2


@dataclass
class RendererSettings:
    max_ray_depth: int = DEFAULT_MAX_RAY_DEPTH
    diffuse_reflection_ray_count: int = DEFAULT_DIFFUSE_REFLECTION_RAY_COUNT
    shadow_bias: float = DEFAULT_SHADOW_BIAS
    reflection_bias: float = DEFAULT_REFLECTION_BIAS
    diffuse_reflection_bias: float = DEFAULT_DIFFUSE_REFLECTION_BIAS
    refraction_bias: float = DEFAULT_REFRACTION_BIAS

    def to_params(self, **overrides) -> RenderParams:
        return default_params(max_ray_depth=self.max_ray_depth, diffuse_reflection_ray_count=self.diffuse_reflection_ray_count,
                              shadow_bias=self.shadow_bias, reflection_bias=self.reflection_bias,
                              diffuse_reflection_bias=self.diffuse_reflection_bias, refraction_bias=self.refraction_bias, **overrides)


def render_image(scene: Scene, settings: RendererSettings | None = None, device: int = 0, **overrides) -> np.ndarray:
    """crt::render_image: float32 [H, W, 3], row 0 = top."""
    r = Renderer(device)
    try:
        r.upload(scene)
        rgb, _, _ = r.render((settings or RendererSettings()).to_params(**overrides), want_rgb8=False)
        return rgb
    finally:
        r.close()


def write_ppm(rgb: np.ndarray, max_color_component: int = 255) -> bytes:
    """Reference text encoder (crt::write_ppm, src/core/crt_image_ppm.cpp:9-23) for small images / tests."""
    h, w, _ = rgb.shape
    v = rgb.astype(np.float32) * np.float32(max_color_component)
    with np.errstate(invalid="ignore"):
        q = np.where(np.isfinite(v) & (np.abs(v) < 2147483648.0), np.trunc(v), -2147483648.0).astype(np.int64)
    q = np.clip(q, 0, max_color_component)
    rows = ["".join(f"{p[0]} {p[1]} {p[2]}\t" for p in row) + "\n" for row in q]
    return (f"P3\n{w} {h}\n{max_color_component}\n" + "".join(rows)).encode()
