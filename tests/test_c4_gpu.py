"""BASELINE config 4 (the bench workload): the seeded 1 002 528-triangle height field, 3840x2160, max_ray_depth 8.
Small instances of the same generator against the oracle on the same inputs; the full-size instance through
size-independent properties (every pipeline bit-identical, k-d walk == reference-order traversal == brute force,
band partition == whole frame) and against the oracle on a low-resolution camera of the full 1 M-triangle scene."""
import numpy as np
import pytest

import oracle_tools as otools

pytestmark = pytest.mark.gpu


def same_bits(a, b):
    return np.array_equal(otools.bits(a), otools.bits(b))


def ok_fraction(a8, b8):
    return float((np.abs(a8.astype(np.int32) - b8.astype(np.int32)).max(axis=2) <= 1).mean())


@pytest.mark.parametrize("material", ["reflective", "diffuse"])
@pytest.mark.parametrize("quads,size", [(48, (256, 144)), (5, (64, 36))])
def test_small_height_fields_against_the_oracle(crt, ot, renderer, material, quads, size):
    from crt_b200 import scenes
    flat = crt.Scene.load_json(scenes.height_field_scene(quads, size[0], size[1], material)).flatten()
    renderer.upload(flat)
    params = crt.default_params(max_ray_depth=8)
    want = ot.port_render(flat, params)
    tri, t, _, _, _ = renderer.primary_hits(crt.TRAVERSAL_KD)
    agree = tri == want["tri"]
    assert agree.mean() >= 0.9999 and same_bits(t[agree], want["t"][agree])
    lit, _, st = renderer.render(crt.default_params(max_ray_depth=8, traversal=crt.TRAVERSAL_LITERAL, collect_stats=1))
    assert same_bits(lit, want["rgb"]) and st["rays"] == want["stats"]["rays"]  # no libm on this path: bit-identical
    for pipeline in (crt.PIPELINE_AUTO, crt.PIPELINE_WAVEFRONT, crt.PIPELINE_HYBRID, crt.PIPELINE_PERSISTENT, crt.PIPELINE_MEGAKERNEL):
        rgb, rgb8, st = renderer.render(crt.default_params(max_ray_depth=8, pipeline=pipeline, collect_stats=1))
        assert ok_fraction(rgb8, want["rgb8"]) >= 0.999
        assert (otools.bits(rgb) == otools.bits(want["rgb"])).all(axis=2).mean() >= 0.9999
        assert abs(st["rays"] - want["stats"]["rays"]) <= 1e-3 * want["stats"]["rays"]


@pytest.fixture(scope="module")
def full_scene(crt):
    from crt_b200 import scenes
    return crt.Scene.load_json(scenes.height_field_scene(708, 3840, 2160, "reflective"))


def test_full_size_c4_properties(crt, ot, full_scene):
    r = crt.Renderer(0)
    try:
        flat = full_scene.flatten()
        assert flat.n_triangles == 1002528 and (flat.camera.width, flat.camera.height) == (3840, 2160)
        r.upload(flat)
        # every pipeline renders the same floats, ray for ray
        frames, rays = {}, {}
        for name, pipeline in (("auto", crt.PIPELINE_AUTO), ("wavefront", crt.PIPELINE_WAVEFRONT), ("hybrid", crt.PIPELINE_HYBRID),
                               ("persistent", crt.PIPELINE_PERSISTENT), ("mega", crt.PIPELINE_MEGAKERNEL)):
            rgb, rgb8, st = r.render(crt.default_params(max_ray_depth=8, pipeline=pipeline, collect_stats=1))
            frames[name], rays[name] = (rgb, rgb8), (st["rays"], st["rays_by_depth"])
        for name in frames:
            assert same_bits(frames[name][0], frames["mega"][0]) and np.array_equal(frames[name][1], frames["mega"][1]), name
            assert rays[name] == rays["mega"], name
        assert rays["mega"][0] > 11_000_000 and rays["mega"][1][8] > 0  # depth 8 is exercised
        # the k-d walk over our SAH tree finds what the reference-order traversal of the reference tree finds
        tri_kd, t_kd, _, _, _ = r.primary_hits(crt.TRAVERSAL_KD, counters=False)
        tri_lit, t_lit, _, _, _ = r.primary_hits(crt.TRAVERSAL_LITERAL, counters=False)
        agree = tri_kd == tri_lit
        assert agree.mean() >= 0.9999, agree.mean()
        assert same_bits(t_kd[agree], t_lit[agree])
        # the PPM text of the frame is what the host encoder makes of the float frame's first rows
        text = r.encode_ppm()
        assert text.startswith(b"P3\n3840 2160\n255\n")
        body = text.split(b"\n", 3)[3]
        first_rows = crt.write_ppm(frames["auto"][0][:2]).split(b"\n", 3)[3]
        assert body.startswith(first_rows)
        # band partition: 8 ranks' bands are the frame
        full = frames["auto"][0]
        for rank in (0, 5):
            part, _, _ = r.render(crt.default_params(max_ray_depth=8, rank=rank, world=8, band_rows=8))
            assert same_bits(part, full[crt.Renderer.owned_rows(2160, rank, 8, 8)])
    finally:
        r.close()


def test_full_scene_low_resolution_against_oracle_and_brute_force(crt, ot, full_scene):
    """same 1 M triangles and tree, 160x90 camera: the oracle and the brute-force kernel finish in seconds"""
    full_scene.set_resolution(160, 90)
    try:
        flat = full_scene.flatten()
        r = crt.Renderer(0)
        try:
            r.upload(flat)
            want = ot.port_render(flat, crt.default_params(max_ray_depth=8))
            tri_kd, t_kd, _, _, _ = r.primary_hits(crt.TRAVERSAL_KD, counters=False)
            tri_b, t_b, _, _, _ = r.primary_hits(crt.TRAVERSAL_BRUTE, counters=False)
            assert (tri_kd == want["tri"]).mean() >= 0.9999
            assert (tri_kd == tri_b).mean() >= 0.9999  # brute force = every triangle in parse order (crt_intersection.cpp:95-107)
            rgb, rgb8, st = r.render(crt.default_params(max_ray_depth=8, collect_stats=1))
            assert ok_fraction(rgb8, want["rgb8"]) >= 0.999 and abs(st["rays"] - want["stats"]["rays"]) <= 1e-3 * want["stats"]["rays"]
        finally:
            r.close()
    finally:
        full_scene.set_resolution(3840, 2160)
