"""crt_cuda_render_frame: every rank's row bands copied to their place in ONE frame in host memory (the frame the
reference's bucket threads assemble in buffer[y*W + x], src/core/crt_renderer.cpp:171-196).  Ranks are played one after
the other on one GPU; ragged heights (a shorter last band, fewer bands than ranks) included."""
import numpy as np
import pytest

import oracle_tools as otools

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("size,world,band", [((160, 90), 3, 8), ((160, 90), 8, 8), ((64, 5), 4, 8), ((96, 64), 2, 16), ((33, 17), 5, 4), ((40, 24), 3, 8)])
def test_bands_of_every_rank_assemble_the_frame_in_host_memory(crt, ot, renderer, size, world, band):
    import os
    rel = "15-01-conclusion/scene1.crtscene"
    scene = crt.Scene.load_json(ot.scene_variant(rel, *size), os.path.dirname(ot.scene_file(rel)))
    renderer.upload(scene.flatten(reference_tree=False))
    W, H = size
    want_rgb, want_rgb8, _ = renderer.render(crt.default_params())
    frame = np.full((H, W, 3), np.nan, np.float32)
    frame8 = np.full((H, W, 3), 77, np.uint8)
    rows_seen = 0
    for rank in range(world):
        params = crt.default_params(rank=rank, world=world, band_rows=band)
        st = renderer.render_frame(params, rgb_frame=frame.ctypes.data, rgb8_frame=frame8.ctypes.data)
        rows_seen += st["rows_rendered"]
    assert rows_seen == H
    assert np.array_equal(otools.bits(frame), otools.bits(want_rgb)) and np.array_equal(frame8, want_rgb8)
    # world <= 1: the same call is crt_cuda_render
    one = np.zeros((H, W, 3), np.float32)
    renderer.render_frame(crt.default_params(), rgb_frame=one.ctypes.data)
    assert np.array_equal(otools.bits(one), otools.bits(want_rgb))


def glass_sheets_scene(width, height, sheets=10):
    """parallel refractive sheets in front of the camera: every hit spawns a reflection that meets the previous sheet and
    a refraction that meets the next one -- the ray tree is binary down to max depth (up to 2^(d+1) - 2 rays per pixel)"""
    import json
    objects = []
    for k in range(1, sheets + 1):
        z = -float(k)
        objects.append({"material_index": 0, "vertices": [-60, -60, z, 60, -60, z, 60, 60, z, -60, 60, z], "triangles": [0, 1, 2, 0, 2, 3]})
    return json.dumps({
        "settings": {"background_color": [0.2, 0.4, 0.8], "image_settings": {"width": width, "height": height, "bucket_size": 8}},
        "camera": {"matrix": [1, 0, 0, 0, 1, 0, 0, 0, 1], "position": [0.013, 0.027, 0.5]},
        "lights": [{"intensity": 100, "position": [0, 5, 0]}],
        "materials": [{"type": "refractive", "ior": 1.5, "smooth_shading": False}],
        "objects": objects})


def test_glass_that_needs_a_hundred_rays_per_pixel_is_rendered_not_dropped(crt, ot):
    """depth 8 in glass needs up to 2^9 - 2 rays per pixel; the wavefront pool starts at one per pixel and
    must be re-sized from the counted demand until the frame fits (or fail loudly) -- never hand back the discarded frame"""
    W, H = 40, 24
    flat = crt.Scene.load_json(glass_sheets_scene(W, H)).flatten()
    r = crt.Renderer(0)  # a fresh context: the pool starts small
    try:
        r.upload(flat)
        params = crt.default_params(max_ray_depth=8, pipeline=crt.PIPELINE_WAVEFRONT, traversal=crt.TRAVERSAL_LITERAL, collect_stats=1)
        want = ot.port_render(flat, params, primary=False)
        assert want["stats"]["rays"] > 100 * W * H, want["stats"]["rays"] / (W * H)  # more than 6 doublings of a one-ray-per-pixel pool
        rgb, rgb8, st = r.render(params)
        assert st["rays"] == want["stats"]["rays"] and st["rays_by_depth"] == want["stats"]["rays_by_depth"]
        assert (np.abs(rgb8.astype(np.int32) - want["rgb8"].astype(np.int32)).max(axis=2) <= 1).mean() >= 0.999
        assert np.nanmax(np.abs(rgb - want["rgb"])) < 2e-3  # powf (Fresnel) is the only libm call
        for pipeline in (crt.PIPELINE_HYBRID, crt.PIPELINE_PERSISTENT, crt.PIPELINE_AUTO):
            again, _, st2 = r.render(crt.default_params(max_ray_depth=8, pipeline=pipeline, collect_stats=1))
            assert st2["rays"] == st["rays"] and np.nanmax(np.abs(again - want["rgb"])) < 2e-3, pipeline
        # and the next scene (no secondary rays) inherits nothing from it
        r.upload(crt.Scene.load_json(ot.scene_variant("14-01-acceleration-tree/scene0.crtscene", W, H)).flatten())
        _, _, st3 = r.render(crt.default_params(collect_stats=1))
        assert st3["rays"] == W * H
    finally:
        r.close()


def test_update_shading_equals_a_fresh_upload(crt, ot, renderer):
    """lights / materials / background / flags changed between two renders must reach the device (the
    reference's render_image reads the Scene on every call); crt_cuda_update_shading re-sends just those"""
    import json
    import os
    rel = "15-01-conclusion/scene1.crtscene"
    root = os.path.dirname(ot.scene_file(rel))
    doc = json.loads(ot.scene_variant(rel, 160, 90))
    base = crt.Scene.load_json(json.dumps(doc), root)
    renderer.upload(base.flatten(reference_tree=False))
    before, _, _ = renderer.render(crt.default_params())
    doc["settings"]["background_color"] = [0.9, 0.1, 0.3]
    doc["settings"]["reflections_on"] = False
    doc["lights"][0]["intensity"] = 2.5 * doc["lights"][0]["intensity"]
    doc["lights"][0]["position"] = [1.0, 6.0, -2.0]
    for m in doc["materials"]:
        if isinstance(m.get("albedo"), list):
            m["albedo"] = [0.3, 0.9, 0.2]
    edited = crt.Scene.load_json(json.dumps(doc), root)
    flat_edited = edited.flatten(reference_tree=False)
    renderer.update_shading(flat_edited)
    updated, _, st = renderer.render(crt.default_params(collect_stats=1))
    assert not np.array_equal(otools.bits(updated), otools.bits(before))
    fresh = crt.Renderer(0)
    try:
        fresh.upload(flat_edited)
        want, _, st_want = fresh.render(crt.default_params(collect_stats=1))
    finally:
        fresh.close()
    assert np.array_equal(otools.bits(updated), otools.bits(want)) and st["rays"] == st_want["rays"]
    # a changed count is refused, not silently mis-rendered
    doc["lights"].append({"intensity": 10, "position": [0, 1, 0]})
    more = crt.Scene.load_json(json.dumps(doc), root).flatten(reference_tree=False)
    with pytest.raises(crt.CrtError):
        renderer.update_shading(more)
