"""csrc/crt_walk_lean.h -- the per-lane k-d walk of the device kernels -- driven one ray at a time on the CPU
(tests/host/walk_lean_host.cpp) against brute force over every triangle with the same triangle test: the walk must
return min (t, parse index) for EVERY ray, i.e. visit every cell that can hold the closest hit.  Camera rays, rays that
start on surfaces (what reflection / refraction / GI rays do), axis-parallel rays that start ON split planes (the
in-plane case: tp = 0 * inf = NaN), rays from outside the root box."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle_tools as otools

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.path.join(HERE, "host", "walk_lean_host.cpp")
LIB = os.path.join(HERE, "host", "_build", "libwalk_lean_host.so")
HEADER = os.path.join(ROOT, "chaos-ray-tracing-course-2025_b200", "csrc", "crt_walk_lean.h")


@pytest.fixture(scope="module")
def host():
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < max(os.path.getmtime(SRC), os.path.getmtime(HEADER)):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math", "-I" + os.path.join(ROOT, "include"), SRC, "-o", LIB],
                       check=True)
    return C.CDLL(LIB)


def closest(host, flat, rays, lazy=0, brute=False):
    rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
    n = len(rays)
    tri, t, visits = np.empty(n, np.int32), np.empty(n, np.float32), np.empty(n, np.int32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    if brute:
        rc = host.brute_closest(C.byref(flat), p(rays), n, p(tri), p(t))
    else:
        rare = C.c_longlong(0)
        rc = host.walk_lean_closest(C.byref(flat), p(rays), n, lazy, p(tri), p(t), p(visits), C.byref(rare))
        closest.rare_events = rare.value
    assert rc == 0, rc
    return tri, t, visits


def same_winners(flat, rays, tri, t, want_tri, want_t, allowed=3):
    """tri / t equal brute force, except for rays that are not a fair question: a ray with a zero direction component
    runs in a plane of constant coordinate, and when that coordinate lies OUTSIDE the bounding box of the triangle brute
    force found (by an ulp or three: the inclusive edge tests of ray_intersect_triangle pass all the same) no spatial
    structure lists the triangle where the ray walks -- the reference's boxes no more than ours."""
    rays = np.asarray(rays, np.float32).reshape(-1, 6)
    verts = flat.arrays()["tri_verts"][:, :, :3]
    bad = np.nonzero((tri != want_tri) | (otools.bits(t) != otools.bits(want_t)))[0]
    unfair = 0
    for i in bad:
        if want_tri[i] < 0:
            return False
        v = verts[want_tri[i]]
        o, d = rays[i, :3], rays[i, 3:]
        outside = any(d[k] == 0 and not (v[:, k].min() <= o[k] <= v[:, k].max()) for k in range(3))
        if not outside:
            return False
        unfair += 1
    return unfair <= allowed


def ray_set(flat, ot, rng, n_primary=6000, n_secondary=6000):
    prim = ot.port_primary_rays(flat).reshape(-1, 6)
    prim = prim[rng.choice(len(prim), min(n_primary, len(prim)), replace=False)]
    lo, hi = np.array(list(flat.root_min), np.float32), np.array(list(flat.root_max), np.float32)
    ext = np.maximum(hi - lo, np.float32(1e-3))
    a = flat.arrays()
    verts = a["tri_verts"][:, :, :3]
    # rays that start on (or a bias away from) surfaces, in every direction
    k = rng.integers(0, len(verts), n_secondary)
    w = rng.dirichlet([1, 1, 1], n_secondary).astype(np.float32)
    origin = (verts[k] * w[:, :, None]).sum(axis=1).astype(np.float32)
    direction = rng.standard_normal((n_secondary, 3)).astype(np.float32)
    direction /= np.linalg.norm(direction, axis=1, keepdims=True)
    origin[: n_secondary // 2] += np.float32(1e-2) * direction[: n_secondary // 2]
    sec = np.concatenate([origin, direction], axis=1)
    # axis-parallel rays starting exactly on split planes of the walk tree (and on vertices: exact coordinates)
    nodes = a["fast_nodes"]  # int32 x 4: {split bits, a, b, axis}
    inner = nodes[nodes[:, 3] < 3]
    special = []
    for _ in range(1500):
        o = (lo + ext * rng.random(3).astype(np.float32)).astype(np.float32)
        on_vertex = False
        if len(inner) and rng.random() < 0.7:
            nd = inner[rng.integers(0, len(inner))]
            o[int(nd[3])] = nd[:1].view(np.float32)[0]
        elif rng.random() < 0.5:
            v = verts[rng.integers(0, len(verts)), rng.integers(0, 3)]
            ax = rng.integers(0, 3)
            o[ax] = v[ax]
            on_vertex = True
        d = np.zeros(3, np.float32)
        kind = rng.random()
        if on_vertex or kind < 0.2:
            # generic direction.  (An axis-parallel ray through a vertex coordinate can run one ulp OUTSIDE a triangle's
            # bounding box and still pass its inclusive edge tests: no spatial structure sees that triangle there, the
            # reference's boxes no more than ours, so such rays are not a fair question.)
            d = rng.standard_normal(3).astype(np.float32)
            d /= np.linalg.norm(d)
        elif kind < 0.6:  # along one axis
            d[rng.integers(0, 3)] = rng.choice([-1.0, 1.0])
        else:  # in a coordinate plane
            i, j = rng.choice(3, 2, replace=False)
            ang = rng.random() * 2 * np.pi
            d[i], d[j] = np.cos(ang), np.sin(ang)
        if rng.random() < 0.3:
            d[d == 0] = -0.0
        if rng.random() < 0.3:  # from outside the root box
            o = (o - d * (ext.max() * np.float32(2))).astype(np.float32)
        special.append(np.concatenate([o, d]))
    return np.concatenate([prim, sec, np.array(special, np.float32)]).astype(np.float32)


SCENES = ["14-01-acceleration-tree/scene1.crtscene", "15-01-conclusion/scene0.crtscene", "15-01-conclusion/scene2.crtscene",
          "11-01-refractive/scene8.crtscene", "11-01-refractive/scene3.crtscene", "09-02-diffuse-smooth-shading/scene3.crtscene",
          "12-01-textures/scene4.crtscene", "14-01-acceleration-tree/scene0.crtscene"]


@pytest.mark.parametrize("rel", SCENES)
def test_lean_walk_finds_the_brute_force_winner_on_bundled_scenes(crt, ot, host, rel):
    flat = crt.Scene.load_json(ot.scene_variant(rel, 160, 90), os.path.dirname(ot.scene_file(rel))).flatten(reference_tree=False)
    rays = ray_set(flat, ot, np.random.default_rng(7))
    want_tri, want_t, _ = closest(host, flat, rays, brute=True)
    for lazy in (0, 3):
        tri, t, visits = closest(host, flat, rays, lazy=lazy)
        assert same_winners(flat, rays, tri, t, want_tri, want_t), (rel, lazy, int((tri != want_tri).sum()))
    assert (want_tri >= 0).mean() > 0.05  # the ray set does hit things


@pytest.mark.parametrize("quads,material", [(48, "reflective"), (5, "diffuse")])
def test_lean_walk_on_height_fields(crt, ot, host, quads, material):
    from crt_b200 import scenes
    flat = crt.Scene.load_json(scenes.height_field_scene(quads, 160, 90, material)).flatten(reference_tree=False)
    rays = ray_set(flat, ot, np.random.default_rng(11))
    want_tri, want_t, _ = closest(host, flat, rays, brute=True)
    tri, t, visits = closest(host, flat, rays, lazy=2)
    assert same_winners(flat, rays, tri, t, want_tri, want_t)


def test_lean_walk_with_a_deep_stack_spills_and_comes_back(crt, ot, host):
    """11-01/scene8's walk tree is 26 levels deep: the 8-entry stack window overflows for some rays, entries go to the
    spill array and come back (kWindowFull / kTopSpilled), and the winners are still the brute-force winners"""
    rel = "11-01-refractive/scene8.crtscene"
    flat = crt.Scene.load_json(ot.scene_variant(rel, 160, 90), os.path.dirname(ot.scene_file(rel))).flatten(reference_tree=False)
    rays = ray_set(flat, ot, np.random.default_rng(3), n_primary=8000, n_secondary=8000)
    want_tri, want_t, _ = closest(host, flat, rays, brute=True)
    tri, t, _ = closest(host, flat, rays, lazy=1)
    assert same_winners(flat, rays, tri, t, want_tri, want_t)
    assert closest.rare_events > 0, "no ray overflowed the stack window: pick a deeper scene"


def test_lean_walk_work_is_bounded(crt, ot, host):
    """the walk is a walk, not a scan: a few dozen node visits per camera ray on the 4 014-triangle scene"""
    rel = "14-01-acceleration-tree/scene1.crtscene"
    flat = crt.Scene.load_json(ot.scene_variant(rel, 160, 90), os.path.dirname(ot.scene_file(rel))).flatten(reference_tree=False)
    rays = ot.port_primary_rays(flat).reshape(-1, 6)
    _, _, visits = closest(host, flat, rays)
    assert visits.mean() < 40 and visits.max() < 2000, (visits.mean(), visits.max())

