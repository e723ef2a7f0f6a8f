"""Host-compiled checks of csrc/field.cuh (it is plain C++ on the host side, where the library's epilogue uses it)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "gnark_icicle_b200", "csrc")


@pytest.mark.parametrize("curve", [254, 377])
def test_binary_inversion_matches_fermat(tmp_path, curve):
    exe = str(tmp_path / f"inv{curve}")
    subprocess.check_call(["g++", "-O2", "-std=c++17", f"-DG16_CURVE={curve}", "-x", "c++", "-I", CSRC,
                           os.path.join(ROOT, "tests", "host", "inv_binary_check.cpp"), "-o", exe])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "0 bad" in out.stdout
