"""The multi-GPU entry points of the C ABI (gblur_mg_create / _run_frames / _run_bands, csrc/multi_gpu.cpp) from a
C++-only program (tests/cpp/mg_check.cpp: no Python, torch or NCCL on the path), checked against the compiled reference.
With one visible GPU the same program runs with n_gpus = 1 (frame blocks and the whole-image fallback of the band call);
with two or more it shards 16384^2 over them."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_mg_capi_from_cpp(gb, tmp_path):
    from libgaussianblur_b200 import build
    from oracle import ref_lib

    exe = tmp_path / "mg_check"
    libdir = os.path.dirname(build.lib_path())
    cmd = ["g++", "-std=c++17", "-O2", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "mg_check.cpp"),
           "-L", libdir, "-lgblur_b200", f"-Wl,-rpath,{libdir}", "-ldl", "-o", str(exe)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    n = min(gb.lib().gblur_device_count(), 8)
    ref = os.path.join(ROOT, "oracle", "_ref", "libgblur_ref.so") if ref_lib.available() else ""
    side = "16384" if n > 1 else "8192"
    r = subprocess.run([str(exe), str(n), ref, side], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "mg check ok" in r.stdout
    print(r.stdout)
