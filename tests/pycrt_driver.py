"""Runs inside a subprocess of tests/test_pycrt_gpu.py: imports ONE `_crt` module (the product's lib/_crt*.so or the
reference's own oracle/_ref/pyref/_crt*.so -- both are called `_crt`, hence one process each) and records what the
Blender add-on would see (reference contract: src/python/py_crt_module.cpp:35-119, 142-157).

    python pycrt_driver.py <module dir> <cases.json> <out.npz>
"""
import json
import sys

import numpy as np


def main():
    module_dir, cases_path, out_path = sys.argv[1:4]
    sys.path.insert(0, module_dir)
    import _crt

    assert _crt.__file__.startswith(module_dir), (_crt.__file__, module_dir)
    cases = json.load(open(cases_path))
    out = {}
    for name, case in cases["renders"].items():
        settings = _crt.RendererSettings(tuple(case["settings"]))
        pixels = _crt.render_scene_from_dict(case["scene"], case["asset_root"], settings)
        assert isinstance(pixels, list), type(pixels)
        assert all(isinstance(p, tuple) and len(p) == 4 for p in pixels[:64])
        assert all(isinstance(v, float) for v in pixels[0])
        out[name] = np.asarray(pixels, dtype=np.float64)
        if hasattr(_crt, "render_scene_from_dict_buffer"):
            w, h, raw = _crt.render_scene_from_dict_buffer(case["scene"], case["asset_root"], settings)
            out[name + "/buffer"] = np.frombuffer(raw, dtype=np.float32).reshape(h, w, 4)
            out[name + "/buffer_wh"] = np.array([w, h])

    # error behaviour, in the order the reference checks: arguments ("O!UO"), scene (ValueError), settings (TypeError)
    good = cases["probe"]  # a loadable scene: the error cases below never reach the render
    good_settings = _crt.RendererSettings(tuple(good["settings"]))
    errors = {}

    def attempt(label, *args):
        try:
            _crt.render_scene_from_dict(*args)
            errors[label] = ["none", ""]
        except Exception as exc:  # noqa: BLE001 - the type and message are what the test compares
            errors[label] = [type(exc).__name__, str(exc)]

    # (a scene whose "settings" lacks "background_color" is undefined behaviour in the reference: its loader reads the
    # member without looking, src/core/crt_json.cpp:550-553, and the module dies with SIGSEGV; not a contract)
    no_materials = dict(good["scene"])
    no_materials.pop("materials", None)  # required and non-empty (src/core/crt_json.cpp:590-592)
    attempt("bad_scene", no_materials, good["asset_root"], good_settings)
    attempt("empty_dict", {}, "/", good_settings)
    attempt("bad_settings", good["scene"], good["asset_root"], (3, 4, 0.01, 0.01, 0.01, 0.01))
    attempt("bad_scene_and_settings", {"nothing": 1}, "/", object())
    attempt("scene_not_a_dict", [1, 2], "/", good_settings)
    attempt("root_not_a_str", good["scene"], b"/", good_settings)
    attempt("too_few_arguments", good["scene"], "/")
    unloadable = dict(good["scene"])
    unloadable.pop("lights", None)  # the loader requires lights (src/core/crt_json.cpp:608-610)
    attempt("scene_without_lights", unloadable, good["asset_root"], good_settings)
    out["errors"] = np.array(json.dumps(errors))

    constants = {k: getattr(_crt, k) for k in dir(_crt) if k.startswith("DEFAULT_")}
    out["constants"] = np.array(json.dumps(constants, sort_keys=True))
    fields = [f for f in ("max_ray_depth", "diffuse_reflection_ray_count", "shadow_bias", "reflection_bias", "diffuse_reflection_bias", "refraction_bias")
              if hasattr(good_settings, f)]
    out["settings_fields"] = np.array(json.dumps({"fields": fields, "len": len(good_settings), "n_fields": _crt.RendererSettings.n_fields,
                                                   "values": [getattr(good_settings, f) for f in fields]}))
    np.savez(out_path, **out)


if __name__ == "__main__":
    main()
