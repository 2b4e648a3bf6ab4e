"""crt_cuda_encode_ppm_noalloc (include/crt_cuda.h): the P3 encoder with its scratch in a static device array and the
text in the caller's device buffer -- no counterpart in the reference.
Checked byte for byte (integer work: the suite's bar is equality) against the existing entry points
crt_cuda_encode_ppm / crt_cuda_encode_ppm_device on the same device frames, several calls in a row on the shared scratch.

Collected last on purpose; passes alone in a fresh process and after the other tests.
"""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _scene(ot, w, h, rel="12-01-textures/scene4.crtscene"):
    return ot.scene_variant(rel, w, h), os.path.dirname(ot.scene_file(rel))


def _device_bytes(n, fill):
    import torch
    t = torch.full((n,), fill, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()  # the library works on its own stream
    return t


def _encode_noalloc(crt, r, rgb8_ptr, w, h, misalign=0):
    """text through the new entry point: length query, then the text into a torch-owned device buffer (the caller's)"""
    length = C.c_size_t()
    rc = crt.lib.crt_cuda_encode_ppm_noalloc(r._ctx, C.c_void_p(rgb8_ptr), w, h, None, 0, C.byref(length))
    assert rc == 0, crt.lib.crt_cuda_last_error(r._ctx)
    pad = 64
    buf = _device_bytes(misalign + length.value + pad, 0xEE)
    again = C.c_size_t()
    rc = crt.lib.crt_cuda_encode_ppm_noalloc(r._ctx, C.c_void_p(rgb8_ptr), w, h, C.c_void_p(buf.data_ptr() + misalign), length.value, C.byref(again))
    assert rc == 0, crt.lib.crt_cuda_last_error(r._ctx)
    assert again.value == length.value
    host = buf.cpu().numpy()
    # nothing outside [misalign, misalign + length) was touched
    assert (host[:misalign] == 0xEE).all() and (host[misalign + length.value:] == 0xEE).all()
    return host[misalign:misalign + length.value].tobytes()


def test_noalloc_ppm_equals_the_existing_entry_points_call_after_call(crt, ot):
    r = crt.Renderer(0)
    try:
        # frames of different sizes one after the other: every call finds the previous call's words in the scratch array
        # (a larger frame first, so that the smaller ones see stale entries beyond their own block count)
        for i, size in enumerate([(640, 360), (37, 19), (1025, 2), (1, 1), (640, 360), (256, 4)]):
            text, root = _scene(ot, *size)
            r.upload(crt.Scene.load_json(text, root))
            rgb, rgb8, _ = r.render(crt.default_params())
            want = r.encode_ppm()                      # existing entry point (ctx-owned, allocated scratch)
            assert want == crt.write_ppm(rgb)          # ... which is the reference's format (crt_image_ppm.cpp:9-23)
            _, rgb8_ptr, w, rows = r.device_framebuffer()
            assert (w, rows) == size
            assert r.encode_ppm_device(rgb8_ptr, w, rows) == want
            first = _encode_noalloc(crt, r, rgb8_ptr, w, rows, misalign=i % 16)
            second = _encode_noalloc(crt, r, rgb8_ptr, w, rows, misalign=(7 * i + 3) % 16)   # two calls in a row
            assert first == want and second == want, size
    finally:
        r.close()


def test_noalloc_ppm_of_a_synthetic_frame_with_every_digit_count(crt):
    """a frame that never went through the renderer: all byte values, ragged width, the caller's own input buffer"""
    import torch
    r = crt.Renderer(0)
    try:
        rng = np.random.default_rng(20251105)
        for (w, h) in [(1031, 97), (3, 5)]:
            frame = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
            frame[0, 0] = (0, 9, 10)
            frame[-1, -1] = (99, 100, 255)
            dev = torch.from_numpy(frame).to("cuda:0")
            torch.cuda.synchronize()
            want = r.encode_ppm_device(dev.data_ptr(), w, h)
            rows = ["".join(f"{p[0]} {p[1]} {p[2]}\t" for p in row) + "\n" for row in frame.tolist()]
            assert want == (f"P3\n{w} {h}\n255\n" + "".join(rows)).encode()
            for k in range(2):
                assert _encode_noalloc(crt, r, dev.data_ptr(), w, h, misalign=5 * k + 1) == want
    finally:
        r.close()


def test_noalloc_ppm_refuses_what_it_cannot_hold(crt):
    r = crt.Renderer(0)
    try:
        frame = _device_bytes(4096, 7)
        length = C.c_size_t(123)
        limit = crt.PPM_NOALLOC_MAX_PIXELS
        # one pixel row beyond the static scratch: refused before anything is launched, with the reason
        rc = crt.lib.crt_cuda_encode_ppm_noalloc(r._ctx, C.c_void_p(frame.data_ptr()), limit + 1, 1, None, 0, C.byref(length))
        assert rc == crt.ERR_INVALID and b"larger than the static scratch" in crt.lib.crt_cuda_last_error(r._ctx)
        rc = crt.lib.crt_cuda_encode_ppm_noalloc(r._ctx, C.c_void_p(frame.data_ptr()), 4096, 2049, None, 0, C.byref(length))
        assert rc == crt.ERR_INVALID
        assert length.value == 123
        # bad arguments / a buffer that is too small
        assert crt.lib.crt_cuda_encode_ppm_noalloc(r._ctx, None, 4, 4, None, 0, C.byref(length)) == crt.ERR_INVALID
        assert crt.lib.crt_cuda_encode_ppm_noalloc(r._ctx, C.c_void_p(frame.data_ptr()), 0, 4, None, 0, C.byref(length)) == crt.ERR_INVALID
        out = _device_bytes(64, 0xEE)
        rc = crt.lib.crt_cuda_encode_ppm_noalloc(r._ctx, C.c_void_p(frame.data_ptr()), 32, 32, C.c_void_p(out.data_ptr()), 64, C.byref(length))
        assert rc == crt.ERR_INVALID and b"too small" in crt.lib.crt_cuda_last_error(r._ctx)
        assert length.value == 13 + 32 * 32 * 6 + 32 and (out.cpu().numpy() == 0xEE).all()   # "P3\n32 32\n255\n" + "7 7 7\t" per pixel + "\n" per row
        # ... and the call still works afterwards
        assert _encode_noalloc(crt, r, frame.data_ptr(), 32, 32) == r.encode_ppm_device(frame.data_ptr(), 32, 32)
    finally:
        r.close()
