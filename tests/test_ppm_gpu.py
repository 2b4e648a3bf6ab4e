"""the P3 text encoder on the device (csrc/crt_ppm.cuh) against crt::write_ppm's format
(src/core/crt_image_ppm.cpp:9-23), byte for byte, through the C ABI and through the crt_renderer CLI."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _scene(ot, w, h, rel="12-01-textures/scene4.crtscene"):
    return ot.scene_variant(rel, w, h), os.path.dirname(ot.scene_file(rel))


# sizes around the encoder's block of 1024 pixels and its 16-byte store alignment, plus ragged and tiny frames
@pytest.mark.parametrize("size", [(1, 1), (37, 19), (1024, 1), (1025, 2), (3, 700), (256, 4), (640, 360)])
def test_device_ppm_equals_reference_format(crt, ot, renderer, size):
    text, root = _scene(ot, *size)
    renderer.upload(crt.Scene.load_json(text, root))
    rgb, rgb8, _ = renderer.render(crt.default_params())
    got = renderer.encode_ppm()
    want = crt.write_ppm(rgb)
    assert got == want
    # and from an arbitrary device pointer (what rank 0 does with the gathered frame)
    _, rgb8_ptr, w, rows = renderer.device_framebuffer()
    assert renderer.encode_ppm_device(rgb8_ptr, w, rows) == want
    # the quantised bytes are the numbers in the text
    body = np.array(got.split(b"\n", 3)[3].split(), dtype=np.int64).reshape(size[1], size[0], 3)
    assert np.array_equal(body, rgb8)


def test_device_ppm_needs_a_frame_and_a_large_enough_buffer(crt, ot):
    import ctypes as C
    r = crt.Renderer(0)
    try:
        with pytest.raises(crt.CrtError):
            r.encode_ppm()
        text, root = _scene(ot, 16, 16)
        r.upload(crt.Scene.load_json(text, root))
        r.render(crt.default_params())
        length = C.c_size_t()
        assert crt.lib.crt_cuda_encode_ppm(r._ctx, None, 0, C.byref(length)) == 0 and length.value > 16 * 16 * 6
        small = C.create_string_buffer(8)
        assert crt.lib.crt_cuda_encode_ppm(r._ctx, small, 8, C.byref(length)) != 0
    finally:
        r.close()


def test_cli_writes_the_same_file_with_either_encoder(crt, ot, tmp_path):
    import __graft_entry__ as entry
    cli = os.path.join(entry.PKG_DIR, "lib", "crt_renderer")
    scene = tmp_path / "s.crtscene"
    text, root = _scene(ot, 320, 180, "15-01-conclusion/scene1.crtscene")
    scene.write_text(text)
    outs = {}
    # group: CRT_DEVICES lists several contexts (here the same GPU twice); literal: the verification traversal, the only
    # mode in which the CLI builds and uploads the reference's tree; lane / packet: both primary-level kernels
    for mode in ("device", "host", "group", "literal", "lane", "packet"):
        out = tmp_path / f"{mode}.ppm"
        env = dict(os.environ, CRT_PPM="host" if mode == "host" else "device")
        if mode == "group":
            env["CRT_DEVICES"] = "0,0"
        if mode == "literal":
            env["CRT_TRAVERSAL"] = "0"
        if mode in ("lane", "packet"):
            env["CRT_PACKET"] = "1" if mode == "packet" else "0"
        proc = subprocess.run([cli, str(scene), str(out)], env=env, capture_output=True, text=True, timeout=120)
        assert proc.returncode == 0 and proc.stdout.startswith("Execution time: "), proc.stderr
        outs[mode] = out.read_bytes()
    assert outs["device"] == outs["host"] == outs["group"] == outs["lane"] == outs["packet"] and outs["device"].startswith(b"P3\n320 180\n255\n")
    a = np.array(outs["device"].split(b"\n", 3)[3].split(), dtype=np.int64)
    b = np.array(outs["literal"].split(b"\n", 3)[3].split(), dtype=np.int64)
    assert a.shape == b.shape and (np.abs(a - b).reshape(-1, 3).max(axis=1) <= 1).mean() >= 0.999
    if ot.have_reference_build():
        ref = tmp_path / "ref.ppm"
        subprocess.run([os.path.join(ot.REF_DIR, "crt_renderer_ref"), str(scene), str(ref)], check=True, capture_output=True, timeout=300)
        a = np.array(outs["device"].split(b"\n", 3)[3].split(), dtype=np.int64)
        b = np.array(ref.read_bytes().split(b"\n", 3)[3].split(), dtype=np.int64)
        assert a.shape == b.shape and (np.abs(a - b).reshape(-1, 3).max(axis=1) <= 1).mean() >= 0.999
        # 15-01/scene1 has no libm call on its path (diffuse + reflective): under the reference-order traversal the CLI's
        # file is the reference CLI's file, byte for byte -- which pins the P3 format of the device encoder to
        # src/core/crt_image_ppm.cpp:9-23 itself rather than to the package's own write_ppm
        assert outs["literal"] == ref.read_bytes()
