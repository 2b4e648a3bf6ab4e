"""The `_crt` CPython entry point (north star: "`_crt` Python/Blender entry points" stay drop-in).

The product's module (chaos-ray-tracing-course-2025_b200/lib/_crt*.so) is compared with the REFERENCE'S OWN module, which
oracle/Makefile builds from /root/reference/src/python/py_crt_module.cpp into oracle/_ref/pyref/.  Both are called
`_crt`, so each is imported in its own subprocess (tests/pycrt_driver.py).

  * not gpu: argument parsing, error types / messages / precedence, module constants, the RendererSettings structseq --
    everything that happens before the render (src/python/py_crt_module.cpp:35-49, 51-53, 88-97, 142-159);
  * gpu: the rendered lists -- W*H 4-tuples, rows bottom-up, alpha 1.0 (:102-116), bit-identical to the reference's
    under the verification traversal on scenes without libm calls, within the bars under the k-d walk; the buffer fast
    path equals the list.
"""
import glob
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import __graft_entry__ as entry

HERE = os.path.dirname(os.path.abspath(__file__))
PRODUCT_DIR = os.path.join(entry.PKG_DIR, "lib")
REFERENCE_DIR = os.path.join(entry.REF_DIR, "pyref")


def _have(module_dir):
    return bool(glob.glob(os.path.join(module_dir, "_crt*.so")))


def _scene_dict(ot, rel, w, h):
    d = json.loads(ot.scene_variant(rel, w, h))
    return d, os.path.dirname(ot.scene_file(rel))


def _run(module_dir, cases, tmp_path, tag, env=None):
    cases_path, out_path = tmp_path / f"{tag}.json", tmp_path / f"{tag}.npz"
    cases_path.write_text(json.dumps(cases))
    proc = subprocess.run([sys.executable, os.path.join(HERE, "pycrt_driver.py"), module_dir, str(cases_path), str(out_path)],
                          env=dict(os.environ, **(env or {})), capture_output=True, text=True, timeout=900)
    assert proc.returncode == 0, proc.stderr[-4000:]
    return dict(np.load(out_path))


def _probe(ot):
    scene, root = _scene_dict(ot, "14-01-acceleration-tree/scene0.crtscene", 32, 18)
    return {"scene": scene, "asset_root": root, "settings": [3, 4, 0.01, 0.01, 0.01, 0.01]}


@pytest.mark.skipif(not (_have(PRODUCT_DIR) and _have(REFERENCE_DIR)), reason="needs both _crt modules (build())")
def test_errors_constants_and_settings_type_equal_the_reference_module(ot, tmp_path):
    cases = {"probe": _probe(ot), "renders": {}}
    ours = _run(PRODUCT_DIR, cases, tmp_path, "ours")
    theirs = _run(REFERENCE_DIR, cases, tmp_path, "theirs")
    e_ours, e_theirs = json.loads(str(ours["errors"])), json.loads(str(theirs["errors"]))
    # what the reference does, stated (so a silently broken reference build cannot make the comparison vacuous) ...
    assert e_theirs["bad_scene"] == ["ValueError", "Invalid CRT Scene dict"]
    assert e_theirs["bad_settings"] == ["TypeError", "Expected a RendererSettings instance"]
    assert e_theirs["bad_scene_and_settings"] == ["ValueError", "Invalid CRT Scene dict"]  # the scene is checked first (:88-97)
    assert e_theirs["scene_without_lights"] == ["ValueError", "Invalid CRT Scene dict"]
    assert e_theirs["scene_not_a_dict"][0] == e_theirs["root_not_a_str"][0] == e_theirs["too_few_arguments"][0] == "TypeError"  # "O!UO"
    # ... and the product does the same, message for message
    assert e_ours == e_theirs
    assert json.loads(str(ours["constants"])) == json.loads(str(theirs["constants"]))
    assert set(json.loads(str(theirs["constants"]))) == {"DEFAULT_SCENE_BUCKET_SIZE", "DEFAULT_MAX_RAY_DEPTH", "DEFAULT_DIFFUSE_REFLECTION_RAY_COUNT",
                                                         "DEFAULT_SHADOW_BIAS", "DEFAULT_REFLECTION_BIAS", "DEFAULT_DIFFUSE_REFLECTION_BIAS",
                                                         "DEFAULT_REFRACTION_BIAS"}
    assert json.loads(str(ours["settings_fields"])) == json.loads(str(theirs["settings_fields"]))
    assert json.loads(str(theirs["settings_fields"]))["n_fields"] == 6


RENDER_CASES = {
    # name: (scene, width, height, settings, libm on the path?)
    "14-01/scene1 defaults": ("14-01-acceleration-tree/scene1.crtscene", 160, 90, [3, 4, 0.01, 0.01, 0.01, 0.01], False),
    "12-01/scene4 bitmap via asset_root": ("12-01-textures/scene4.crtscene", 160, 90, [3, 4, 0.01, 0.01, 0.01, 0.01], False),
    "15-01/scene1 reflective depth 5 bias 0.02": ("15-01-conclusion/scene1.crtscene", 128, 72, [5, 4, 0.01, 0.02, 0.01, 0.01], False),
    "11-01/scene8 refractive defaults": ("11-01-refractive/scene8.crtscene", 160, 90, [3, 4, 0.01, 0.01, 0.01, 0.01], True),
    "11-01/scene8 refractive depth 6": ("11-01-refractive/scene8.crtscene", 96, 54, [6, 2, 0.02, 0.005, 0.01, 0.03], True),
    "15-01/scene2 GI 2 rays": ("15-01-conclusion/scene2.crtscene", 48, 48, [2, 2, 0.01, 0.01, 0.01, 0.01], True),
    "ragged 37x19": ("14-01-acceleration-tree/scene0.crtscene", 37, 19, [3, 4, 0.01, 0.01, 0.01, 0.01], False),
}


@pytest.mark.gpu
@pytest.mark.skipif(not (_have(PRODUCT_DIR) and _have(REFERENCE_DIR)), reason="needs both _crt modules (build())")
def test_rendered_lists_equal_the_reference_module(ot, tmp_path):
    cases = {"probe": _probe(ot), "renders": {}}
    for name, (rel, w, h, settings, _) in RENDER_CASES.items():
        scene, root = _scene_dict(ot, rel, w, h)
        cases["renders"][name] = {"scene": scene, "asset_root": root, "settings": settings}
    theirs = _run(REFERENCE_DIR, cases, tmp_path, "theirs")
    literal = _run(PRODUCT_DIR, cases, tmp_path, "ours_literal", env={"CRT_TRAVERSAL": "0"})  # reference-order traversal
    walk = _run(PRODUCT_DIR, cases, tmp_path, "ours_kd")                                       # the product's default
    for name, (rel, w, h, settings, libm) in RENDER_CASES.items():
        want = theirs[name]
        assert want.shape == (w * h, 4) and (want[:, 3] == 1.0).all()
        for got in (literal[name], walk[name]):
            assert got.shape == want.shape and got.dtype == want.dtype and (got[:, 3] == 1.0).all(), name
        # rows bottom-up: list row 0 is the image's LAST row.  Checked against the C ABI's top-down frame below via
        # the reference itself: equal lists mean equal orientation; and the two orientations differ on these scenes
        img = want.reshape(h, w, 4)
        assert not np.array_equal(img, img[::-1]), name
        if not libm:  # no powf / sinf / cosf on the path: the floats are the reference's, bit for bit
            assert np.array_equal(literal[name].view(np.uint64), want.view(np.uint64)), name
        for got in (literal[name], walk[name]):  # the PPM bar on what write_ppm would make of it (crt_image_ppm.cpp:16-19)
            q = lambda a: np.clip((a[:, :3].astype(np.float32) * np.float32(255)).astype(np.int64), 0, 255)
            assert (np.abs(q(got) - q(want)).max(axis=1) <= 1).mean() >= 0.999, name
        # the buffer fast path is the list, reshaped (float32 RGBA, rows bottom-up)
        buf = walk[name + "/buffer"]
        assert tuple(walk[name + "/buffer_wh"]) == (w, h) and buf.shape == (h, w, 4)
        assert np.array_equal(buf.reshape(-1, 4).astype(np.float64), walk[name]), name
