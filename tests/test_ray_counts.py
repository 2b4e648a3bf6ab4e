"""tests/golden/ray_counts.json -- the rays per frame bench.py's reference arm divides by (the reference cannot count its
own rays, and that arm must not load the product).  Written by tools/count_rays.py from the device's counters; pinned
here to the oracle port on the CPU (same recursion as the reference, bit for bit) and, in the GPU suite, to the
device's counters on every workload."""
import json
import os

import numpy as np
import pytest

import bench

HERE = os.path.dirname(os.path.abspath(__file__))
TABLE = json.load(open(os.path.join(HERE, "golden", "ray_counts.json")))


def test_table_covers_every_bench_workload():
    assert set(TABLE) == set(bench.WORKLOADS)
    for name, (desc, W, H, depth, _) in bench.WORKLOADS.items():
        e = TABLE[name]
        assert (e["width"], e["height"], e["max_ray_depth"]) == (W, H, depth) and e["rays_per_frame"] >= W * H
    assert len(TABLE["c5"]["rays_per_orbit_frame"]) == 240


def _port_rays(crt, ot, text, root, depth):
    flat = crt.Scene.load_json(text, root).flatten()
    return ot.port_render(flat, crt.default_params(max_ray_depth=depth, traversal=crt.TRAVERSAL_LITERAL), primary=False)["stats"]


def test_c1_and_c3_counts_equal_the_oracle_ports(crt, ot):
    for name in ("c1", "c3"):
        text, root = bench.scene_text(name)
        st = _port_rays(crt, ot, text, root, bench.WORKLOADS[name][3])
        assert st["rays"] == TABLE[name]["rays_per_frame"], name
        assert [v for v in st["rays_by_depth"] if v] == TABLE[name]["rays_by_depth"], name


@pytest.mark.parametrize("k", [0, 97, 239])
def test_c5_orbit_frame_counts_equal_the_oracle_ports(crt, ot, k):
    text, root = bench.scene_text("c5")
    pos, mat = bench.orbit_of(text)[k]
    doc = json.loads(text)
    doc["camera"]["position"], doc["camera"]["matrix"] = [float(v) for v in pos], [float(v) for v in mat]
    st = _port_rays(crt, ot, json.dumps(doc), root, 3)
    assert st["rays"] == TABLE["c5"]["rays_per_orbit_frame"][k]


def test_orbit_from_the_scene_text_is_the_orbit_from_the_flattened_scene(crt, ot):
    """bench.orbit_of (no product loaded: the reference arm) against scenes.orbit_cameras over the flattener's root box"""
    from crt_b200 import scenes
    text, root = bench.scene_text("c5")
    flat = crt.Scene.load_json(text, root).flatten(reference_tree=False)
    center = (np.array(list(flat.root_min)) + np.array(list(flat.root_max))) / 2
    want = scenes.orbit_cameras(list(flat.camera.position), list(flat.camera.rotation), center, 240)
    got = bench.orbit_of(text)
    for (p0, m0), (p1, m1) in zip(want, got):
        assert np.array_equal(p0, p1) and np.array_equal(m0, m1)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(bench.WORKLOADS))
def test_device_counts_equal_the_table(crt, ot, name):
    desc, W, H, depth, _ = bench.WORKLOADS[name]
    text, root = bench.scene_text(name)
    flat = crt.Scene.load_json(text, root).flatten(reference_tree=False)
    r = crt.Renderer(0)
    try:
        r.upload(flat)
        params = crt.default_params(max_ray_depth=depth, collect_stats=1)
        if name == "c5":
            orbit = bench.orbit_of(text)
            for k in (0, 1, 60, 137, 239):
                cam = crt.Camera.from_buffer_copy(flat.camera)
                cam.position[:] = [float(v) for v in orbit[k][0]]
                cam.rotation[:] = [float(v) for v in orbit[k][1]]
                r.set_camera(cam)
                _, _, st = r.render(params, want_rgb=False, want_rgb8=False)
                assert st["rays"] == TABLE[name]["rays_per_orbit_frame"][k], k
        else:
            _, _, st = r.render(params, want_rgb=False, want_rgb8=False)
            assert st["rays"] == TABLE[name]["rays_per_frame"]
            assert [v for v in st["rays_by_depth"] if v] == TABLE[name]["rays_by_depth"]
            assert 0 < st["entered_rays"] <= st["rays"]
    finally:
        r.close()
