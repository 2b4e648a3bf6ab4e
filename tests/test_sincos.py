"""The GI bounce angles' sine / cosine as the device computes them (csrc/crt_sincos.h, compiled here for the host) over
every argument the renderer can produce: equal to binary64 sin / cos rounded once to binary32, the value the device has
always used and DESIGN.md documents as the FP difference from glibc's sinf / cosf (src/core/crt_renderer.cpp:68-74)."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_device_sincos_equals_rounded_binary64_on_every_possible_argument(tmp_path):
    exe = str(tmp_path / "sincos_check")
    subprocess.run(["g++", "-O2", "-ffp-contract=off", os.path.join(ROOT, "tools", "sincos_check.cpp"), "-o", exe], check=True)
    out = subprocess.run([exe], stdout=subprocess.PIPE, text=True, check=True).stdout
    assert re.search(r"arguments 16777216\b", out)
    assert "sincos_0_2pi vs rounded binary64: sin 0 cos 0" in out
    sin_diff, cos_diff = (int(v) for v in re.search(r"glibc sinf/cosf vs rounded binary64: sin (\d+) cos (\d+)", out).groups())
    assert 0 < sin_diff < 16777216 * 0.02 and 0 < cos_diff < 16777216 * 0.02, "the documented last-bit difference: ~1.3 % of the arguments"
