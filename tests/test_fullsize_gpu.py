"""BASELINE.json configs at their STATED sizes against the LIVE reference (oracle/_ref/ref_harness = the unmodified
reference's read_scene_from_istream + render_image + ray_intersect_acceleration_tree, src/core/crt_renderer.cpp:147-199,
crt_intersection.cpp:109-136), once per suite:

  C4  the 1 002 528-triangle height field, 3840x2160, max_ray_depth 8          (configs[3], the bench workload)
  C3  15-01-conclusion/scene0,1 and 12-01-textures/scene4 at 3840x2160;         (configs[2])
      15-01/scene2 with GI at 540x540 against the reference itself, not the port
  C5  five frames of the 240-frame orbit of 11-01-refractive/scene8, 1920x1080, (configs[4])
      the reference rendering the same rewritten JSON (scenes.orbit_cameras: float64 look-at, rounded once)

Bars (BASELINE.json): same primary hit triangle on >= 99.99 % of rays, PPM channels within 1/255 on >= 99.9 % of
pixels; bit-identical floats under the verification traversal where no libm call is on the path.
"""
import json
import os

import numpy as np
import pytest

import oracle_tools as otools

pytestmark = pytest.mark.gpu


def quantised(rgb):
    return np.clip(np.trunc(np.nan_to_num(rgb.astype(np.float32) * np.float32(255), nan=-1.0)), 0, 255).astype(np.int32)


def ppm_ok_fraction(rgb_a, rgb_b):
    return float((np.abs(quantised(rgb_a) - quantised(rgb_b)).max(axis=2) <= 1).mean())


def same_bits(a, b):
    return np.array_equal(otools.bits(a), otools.bits(b))


def needs_reference(ot):
    if not ot.have_reference_build():
        pytest.skip("oracle/_ref (the compiled reference) did not travel")


def test_c4_at_3840x2160_depth_8_against_the_live_reference(crt, ot, tmp_path):
    needs_reference(ot)
    from crt_b200 import scenes
    text = scenes.height_field_scene(708, 3840, 2160, "reflective")
    path = tmp_path / "c4.crtscene"
    path.write_text(text)
    ref = ot.run_harness(str(path), render=True, hits=True, max_depth=8)
    assert ref.get("loaded") and (ref["width"], ref["height"]) == (3840, 2160), {k: v for k, v in ref.items() if not hasattr(v, "shape")}
    scene = crt.Scene.load_json(text)
    r = crt.Renderer(0)
    try:
        flat = scene.flatten()  # with the reference's tree: the verification traversal below needs it
        assert flat.n_triangles == 1002528
        r.upload(flat)
        tri, t, _, _, _ = r.primary_hits(crt.TRAVERSAL_KD, counters=False)
        agree = tri == ref["tri"]
        assert agree.mean() >= 0.9999, agree.mean()
        assert same_bits(t[agree & (tri >= 0)], ref["t"][agree & (tri >= 0)])
        rgb, rgb8, stats = r.render(crt.default_params(max_ray_depth=8, collect_stats=1))
        assert ppm_ok_fraction(rgb, ref["rgb"]) >= 0.999
        assert np.array_equal(quantised(rgb), rgb8.astype(np.int32))
        identical = (otools.bits(rgb) == otools.bits(ref["rgb"])).all(axis=2).mean()
        assert identical >= 0.9999, identical
        # the committed ray count the reference arm of bench.py divides by (it cannot count rays itself)
        counts = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "ray_counts.json")))
        assert stats["rays"] == counts["c4"]["rays_per_frame"], stats["rays"]
        # reference-order traversal of the reference's own tree: the floats are the reference's (no libm on this path)
        lit, _, _ = r.render(crt.default_params(max_ray_depth=8, traversal=crt.TRAVERSAL_LITERAL))
        assert same_bits(lit, ref["rgb"])
    finally:
        r.close()


@pytest.mark.parametrize("rel", ["15-01-conclusion/scene0.crtscene", "15-01-conclusion/scene1.crtscene", "12-01-textures/scene4.crtscene"])
def test_c3_at_3840x2160_against_the_live_reference(crt, ot, renderer, rel):
    needs_reference(ot)
    text, root = ot.scene_variant(rel, 3840, 2160), os.path.dirname(ot.scene_file(rel))
    ref = ot.run_harness(text, asset_root=root, render=True, hits=True)
    assert ref.get("loaded") and (ref["width"], ref["height"]) == (3840, 2160)
    flat = crt.Scene.load_json(text, root).flatten()
    renderer.upload(flat)
    tri, t, _, _, _ = renderer.primary_hits(crt.TRAVERSAL_KD, counters=False)
    agree = tri == ref["tri"]
    assert agree.mean() >= 0.9999, agree.mean()
    assert same_bits(t[agree & (tri >= 0)], ref["t"][agree & (tri >= 0)])
    rgb, rgb8, _ = renderer.render(crt.default_params())
    assert ppm_ok_fraction(rgb, ref["rgb"]) >= 0.999
    lit, _, _ = renderer.render(crt.default_params(traversal=crt.TRAVERSAL_LITERAL))
    assert same_bits(lit, ref["rgb"])  # diffuse / reflective / textures: no libm call


def test_c3_scene2_with_gi_at_540x540_against_the_live_reference(crt, ot, renderer):
    needs_reference(ot)
    rel = "15-01-conclusion/scene2.crtscene"
    text, root = ot.scene_variant(rel, 540, 540), os.path.dirname(ot.scene_file(rel))
    ref = ot.run_harness(text, asset_root=root, render=True, hits=True)
    assert ref.get("loaded") and (ref["width"], ref["height"]) == (540, 540)
    flat = crt.Scene.load_json(text, root).flatten()
    assert flat.gi_on
    renderer.upload(flat)
    tri, _, _, _, _ = renderer.primary_hits(crt.TRAVERSAL_KD, counters=False)
    assert (tri == ref["tri"]).mean() >= 0.9999
    for traversal in (crt.TRAVERSAL_KD, crt.TRAVERSAL_LITERAL):
        rgb, _, _ = renderer.render(crt.default_params(traversal=traversal))
        assert ppm_ok_fraction(rgb, ref["rgb"]) >= 0.999, traversal
        # sinf / cosf (GI directions) and powf (Fresnel) are glibc's on the CPU: last-bit differences, far below 1/255
        assert np.nanmax(np.abs(rgb - ref["rgb"])) < 2e-3, traversal


def test_c5_orbit_frames_against_the_live_reference(crt, ot, renderer):
    needs_reference(ot)
    from crt_b200 import scenes
    rel = "11-01-refractive/scene8.crtscene"
    root = os.path.dirname(ot.scene_file(rel))
    base = json.loads(ot.scene_variant(rel, 1920, 1080))
    flat0 = crt.Scene.load_json(json.dumps(base), root).flatten(reference_tree=False)
    center = (np.array(list(flat0.root_min)) + np.array(list(flat0.root_max))) / 2
    orbit = scenes.orbit_cameras(list(flat0.camera.position), list(flat0.camera.rotation), center, 240)
    renderer.upload(flat0)
    for k in (0, 1, 60, 137, 239):
        pos, mat = orbit[k]
        doc = json.loads(json.dumps(base))
        doc["camera"]["position"] = [float(v) for v in pos]   # binary32 values: repr of the widened double round-trips
        doc["camera"]["matrix"] = [float(v) for v in mat]
        text = json.dumps(doc)
        ref = ot.run_harness(text, asset_root=root, render=True, hits=True)
        assert ref.get("loaded")
        # the way bench.py renders the orbit: one upload, a new 72-byte camera per frame
        cam = crt.Camera.from_buffer_copy(flat0.camera)
        cam.position[:] = [float(v) for v in pos]
        cam.rotation[:] = [float(v) for v in mat]
        renderer.set_camera(cam)
        tri, _, _, _, _ = renderer.primary_hits(crt.TRAVERSAL_KD, counters=False)
        assert (tri == ref["tri"]).mean() >= 0.9999, k
        rgb, _, _ = renderer.render(crt.default_params())
        assert ppm_ok_fraction(rgb, ref["rgb"]) >= 0.999, k
        assert np.nanmax(np.abs(rgb - ref["rgb"])) < 2e-3, k  # powf (Fresnel) is the only libm call on this path
        # ... and the same frame from the rewritten JSON through the loader
        flat_k = crt.Scene.load_json(text, root).flatten(reference_tree=False)
        assert bytes(flat_k.camera) == bytes(cam), k
    renderer.set_camera(crt.Camera.from_buffer_copy(flat0.camera))
