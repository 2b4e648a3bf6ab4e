"""Deterministic synthetic workloads (written as ordinary .crtscene JSON so the reference and the GPU path read the
same file).  Definitions follow SURVEY.md §8(d)."""
from __future__ import annotations

import json
import math

import numpy as np


def _join(values, fmt="%.9g"):
    # %.9g round-trips every binary32 value
    return ",".join(fmt % v for v in values)


def height_field_scene(quads: int = 708, width: int = 3840, height: int = 2160, material: str = "reflective", seed: int = 1234) -> str:
    """C4: seeded height field of quads x quads cells = 2*quads^2 triangles (708 -> 1 002 528), one object, one point
    light, camera at (0, 14, 14) pitched 45 degrees down, fov 90."""
    n = quads + 1
    rng = np.random.default_rng(seed)
    axis = np.linspace(-10.0, 10.0, n, dtype=np.float32)
    x, z = np.meshgrid(axis, axis)  # row i = z index, column j = x index
    y = (np.float32(0.6) * np.sin(np.float32(1.3) * x) * np.cos(np.float32(1.7) * z)).astype(np.float32)
    y = (y + np.float32(0.05) * rng.standard_normal((n, n), dtype=np.float32)).astype(np.float32)
    verts = np.stack([x, y, z], axis=-1).reshape(-1)
    i, j = np.meshgrid(np.arange(quads), np.arange(quads), indexing="ij")
    a = (i * n + j).reshape(-1)
    b, c, d = a + 1, a + n, a + n + 1
    tris = np.stack([a, c, b, b, c, d], axis=1).reshape(-1)
    s = math.sqrt(0.5)
    mat = {"type": material, "albedo": [0.8, 0.8, 0.8], "smooth_shading": True}
    head = {
        "settings": {"background_color": [0.1, 0.2, 0.4], "image_settings": {"width": width, "height": height, "bucket_size": 24}},
        "camera": {"matrix": [1, 0, 0, 0, float(np.float32(s)), -float(np.float32(s)), 0, float(np.float32(s)), float(np.float32(s))],
                   "position": [0, 14, 14], "fov_degrees": 90},
        "lights": [{"intensity": 3000, "position": [0, 12, 0]}],
        "materials": [mat],
    }
    text = json.dumps(head)[:-1]
    text += ',"objects":[{"material_index":0,"vertices":[' + _join(verts.tolist()) + '],"triangles":[' + ",".join(map(str, tris.tolist())) + "]}]}"
    return text


def look_at(position, target, up=(0.0, 1.0, 0.0)):
    """Row-major camera matrix whose rows are the camera's right / up / back axes (rays are row-vector * matrix and the
    camera looks down its local -z, src/core/crt_camera.cpp:28-31).  float64 maths, rounded once by the caller."""
    p, t, u = (np.asarray(v, dtype=np.float64) for v in (position, target, up))
    back = p - t
    back /= np.linalg.norm(back)
    right = np.cross(u, back)
    right /= np.linalg.norm(right)
    true_up = np.cross(back, right)
    return np.stack([right, true_up, back])


def orbit_cameras(position, matrix, center, frames: int = 240):
    """C5: camera on a circle about `center` at the original radius and height, yaw = 2*pi*k/frames, looking at the
    centre's height-projected point the original camera looked towards.  Returns [(position[3], matrix[9])]."""
    p0 = np.asarray(position, dtype=np.float64)
    c = np.asarray(center, dtype=np.float64)
    radius = math.hypot(p0[0] - c[0], p0[2] - c[2])
    phase = math.atan2(p0[0] - c[0], p0[2] - c[2])
    out = []
    for k in range(frames):
        a = phase + 2.0 * math.pi * k / frames
        pos = np.array([c[0] + radius * math.sin(a), p0[1], c[2] + radius * math.cos(a)])
        m = look_at(pos, c)
        out.append((pos.astype(np.float32), m.astype(np.float32).reshape(-1)))
    return out
