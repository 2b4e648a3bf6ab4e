"""The C-ABI library loads and exports every symbol include/crt_cuda.h declares; struct layouts match the header.
No compute calls here (no GPU)."""
import ctypes
import os
import re
import subprocess
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "crt_cuda.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(crt_(?:cuda|scene|render)_\w+)\s*\(", text)))


def test_every_declared_symbol_is_exported(crt):
    names = declared_functions()
    assert len(names) >= 20
    for name in names:
        assert hasattr(crt.lib, name), f"{name} declared in include/crt_cuda.h but not exported by libcrt_b200.so"
    assert sorted(crt.EXPORTED_SYMBOLS) == names


def test_struct_layouts_match_header(crt):
    prog = r'''
    #include <stdio.h>
    #include "crt_cuda.h"
    int main(void) {
        printf("%zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(crt_box_node), sizeof(crt_kd_node), sizeof(crt_material), sizeof(crt_texture),
               sizeof(crt_camera), sizeof(crt_flat_scene), sizeof(crt_render_params), sizeof(crt_render_stats));
        return 0;
    }'''
    with tempfile.TemporaryDirectory() as tmp:
        src = os.path.join(tmp, "s.c")
        open(src, "w").write(prog)
        exe = os.path.join(tmp, "s")
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), src, "-o", exe], check=True)
        sizes = [int(v) for v in subprocess.run([exe], stdout=subprocess.PIPE, text=True, check=True).stdout.split()]
    assert sizes[:4] == [32, 16, 16, 48]
    assert sizes[4] == ctypes.sizeof(crt.Camera)
    assert sizes[5] == ctypes.sizeof(crt.FlatScene)
    assert sizes[6] == ctypes.sizeof(crt.RenderParams)
    assert sizes[7] == ctypes.sizeof(crt.RenderStats)


def test_defaults_mirror_reference_settings(crt):
    # src/core/crt_renderer.h:10-25
    p = crt.default_params()
    assert (p.max_ray_depth, p.diffuse_reflection_ray_count) == (3, 4)
    for f in ("shadow_bias", "reflection_bias", "diffuse_reflection_bias", "refraction_bias"):
        assert getattr(p, f) == pytest.approx(1e-2)
    assert p.shadows == 0 and p.world == 1
    s = crt.RendererSettings()
    assert (s.max_ray_depth, s.diffuse_reflection_ray_count, s.shadow_bias) == (3, 4, 1e-2)


def test_no_device_is_a_loud_error_not_a_fallback(crt):
    if crt.lib.crt_cuda_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(crt.CrtError):
        crt.Renderer(0)


def test_null_arguments_are_rejected(crt):
    assert crt.lib.crt_scene_load_file(None, None) == crt.ERR_INVALID
    assert crt.lib.crt_cuda_create(0, None) == crt.ERR_INVALID
    assert crt.lib.crt_cuda_render(None, None, None, None, None) == crt.ERR_INVALID
    h = ctypes.c_void_p()
    assert crt.lib.crt_scene_load_file(b"/nonexistent/scene.crtscene", ctypes.byref(h)) == crt.ERR_IO
