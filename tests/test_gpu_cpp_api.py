"""The reference's gtest intents re-asserted against the drop-in C++ API: compiles tests/cpp/api_intents.cpp with g++
against include/ and libgblur_b200.so and runs it on the GPU. Also the reference's Python stress intent
(python/tests/stress.py:23-47: 10240^2 RGBA, sigma in {25, 50, 100, 150}, each < 10 s, variance decreases) through the
pybind11 module's zero-copy entry point."""
import os
import subprocess
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_gtest_intents_through_cpp_api(gb, tmp_path):
    from libgaussianblur_b200 import build

    exe = tmp_path / "api_intents"
    libdir = os.path.dirname(build.lib_path())
    cmd = ["g++", "-std=c++17", "-O2", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "api_intents.cpp"),
           "-pthread", "-L", libdir, "-lgblur_b200", f"-Wl,-rpath,{libdir}", "-o", str(exe)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "api intents ok" in r.stdout
    assert "Unsupported number of channels" in r.stderr and "Invalid smoothing factor" in r.stdout


def test_python_stress_intent_through_pybind_module(gb):
    m = gb.gaussianblur_module()
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, size=(10240, 10240, 4), dtype=np.uint8)
    times = []
    for sigma in (25.0, 50.0, 100.0, 150.0):
        before = float(img[::16, ::16].astype(np.float64).var())
        t0 = time.perf_counter()
        m.gaussianblur_array(img, sigma, True)
        times.append(time.perf_counter() - t0)
        assert float(img[::16, ::16].astype(np.float64).var()) < before
    assert max(times) < 10.0 and max(times[1:]) - min(times[1:]) < 2.0, times


def test_pybind_module_list_surface(gb):
    """The reference's own Python usage (python/tests/stress.py:11-29): Image.data as a list."""
    m = gb.gaussianblur_module()
    from oracle import gblur_oracle as go

    src = go.synth_u8((24, 40, 4), 3)
    image = m.Image()
    image.geom = m.ImgGeom()
    image.geom.rows, image.geom.cols, image.geom.channels = 24, 40, 4
    g = m.ImgGeom()
    g.rows, g.cols, g.channels = 24, 40, 4
    image.geom = g
    image.data = src.reshape(-1).tolist()
    m.gaussianblur(image=image, sigma=2.0, apply_to_alpha=False)
    out = np.array(image.data, dtype=np.uint8).reshape(24, 40, 4)
    want = go.gaussianblur(src, 2.0, False)
    assert np.abs(out.astype(int) - want.astype(int)).max() <= 1
    assert np.array_equal(out[..., 3], src[..., 3])
