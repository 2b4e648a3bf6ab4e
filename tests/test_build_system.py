"""Build-system drop-in: the top-level CMakeLists.txt keeps the reference's targets `crt_core`,
`crt_renderer`, `_crt` and its options, adds the CUDA language for sm_100a; the Makefile keeps `standalone` / `python` /
`clean` (/root/reference/CMakeLists.txt:24-58, /root/reference/Makefile:31-59)."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cmake_lists_declares_the_reference_targets_and_sm_100a():
    text = open(os.path.join(ROOT, "CMakeLists.txt")).read()
    assert re.search(r"project\(crt_renderer LANGUAGES CXX CUDA\)", text)
    assert re.search(r"add_library\(crt_core STATIC", text) and re.search(r"add_executable\(\$\{PROJECT_NAME\}", text)
    assert re.search(r"Python3_add_library\(_crt MODULE WITH_SOABI", text)
    assert "CMAKE_CUDA_ARCHITECTURES 100a" in text and "option(BUILD_STANDALONE" in text and "option(BUILD_PYTHON" in text
    make = open(os.path.join(ROOT, "Makefile")).read()
    for target in ("standalone:", "python:", "clean:"):
        assert target in make


@pytest.mark.skipif(not (shutil.which("cmake") and os.path.exists("/usr/local/cuda/bin/nvcc")), reason="needs cmake and nvcc")
def test_make_standalone_builds_crt_renderer_for_sm_100a():
    proc = subprocess.run(["make", "standalone"], cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
    exe = os.path.join(ROOT, "build", "standalone", "crt_renderer")
    assert os.path.exists(exe)
    cubins = subprocess.run(["cuobjdump", "-lelf", exe], capture_output=True, text=True).stdout
    assert "sm_100a" in cubins and not re.search(r"sm_(?!100a)\d+", cubins), cubins
    # the CLI contract holds for this binary too (src/standalone/main.cpp:17-20)
    proc = subprocess.run([exe, "/nonexistent.crtscene", "/tmp/never.ppm"], capture_output=True, text=True, timeout=60)
    assert proc.returncode == 1 and proc.stderr.startswith("Error: Could not open input file: ")
