"""The crt_renderer CLI's contract where no device is needed (src/standalone/main.cpp:11-48): messages, exit codes,
and -- on a box without a GPU -- a loud failure instead of a CPU fallback."""
import os
import subprocess

import pytest


def _scene(ot, w, h, rel="14-01-acceleration-tree/scene0.crtscene"):
    return ot.scene_variant(rel, w, h), os.path.dirname(ot.scene_file(rel))


def test_cli_error_messages_and_exit_codes_are_the_reference_clis(crt, ot, tmp_path):
    """src/standalone/main.cpp:17-33: three messages on stderr, exit code 1"""
    import __graft_entry__ as entry
    cli = os.path.join(entry.PKG_DIR, "lib", "crt_renderer")
    bad = tmp_path / "bad.crtscene"
    bad.write_text('{"settings": 1}')
    good = tmp_path / "good.crtscene"
    good.write_text(_scene(ot, 16, 9)[0])
    cases = [([str(tmp_path / "missing.crtscene"), str(tmp_path / "o.ppm")], "Error: Could not open input file: "),
             ([str(bad), str(tmp_path / "o.ppm")], "Error: Could not parse JSON file: "),
             ([str(good), str(tmp_path / "no_such_dir" / "o.ppm")], "Error: Could not open output file: ")]
    binaries = [cli] + ([os.path.join(ot.REF_DIR, "crt_renderer_ref")] if ot.have_reference_build() else [])
    for argv, message in cases:
        seen = []
        for binary in binaries:
            proc = subprocess.run([binary] + argv, capture_output=True, text=True, timeout=120)
            assert proc.returncode == 1 and proc.stderr.startswith(message), (binary, argv, proc.returncode, proc.stderr)
            seen.append(proc.stderr)
        assert len(set(seen)) == 1, seen


def test_cli_fails_loudly_without_a_device(crt, ot, tmp_path):
    """no CUDA device: exit code 1 and a message, never a silently rendered frame (there is no CPU path)"""
    import __graft_entry__ as entry
    if crt.lib.crt_cuda_device_count() > 0:
        pytest.skip("this box has a GPU")
    cli = os.path.join(entry.PKG_DIR, "lib", "crt_renderer")
    good = tmp_path / "good.crtscene"
    good.write_text(_scene(ot, 16, 9)[0])
    out = tmp_path / "o.ppm"
    proc = subprocess.run([cli, str(good), str(out)], capture_output=True, text=True, timeout=120)
    assert proc.returncode == 1 and "GPU only" in proc.stderr and "Execution time" not in proc.stdout
