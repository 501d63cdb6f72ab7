"""Overlap-save row pass (csrc/ols_core.cuh, csrc/ols_kernels.cu), checked WITHOUT a GPU: the
kernel's lane functions are plain C++ as well, so tests/cpp/ols_emulate.cpp runs them lane by lane on the host (same
tables, same panel layout, same index maps) and this file compares the result with the float64 oracle of the
reference's row pass (process_channel_tiles, /root/reference/src/gaussianblur.cpp:202-244, including the Nyquist
behaviour of :62-79) and, where it was built, with the compiled reference itself.
Tolerance: 5e-4 absolute on the fp32 intermediate (the bar tests/test_gpu_parity.py uses for the row pass)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import gblur_oracle as go
from oracle import ref_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 5e-4


@pytest.fixture(scope="module")
def emu(tmp_path_factory):
    so = tmp_path_factory.mktemp("ols") / "libols_emu.so"
    cmd = ["g++", "-std=c++17", "-O2", "-shared", "-fPIC", "-w", "-I", "/usr/local/cuda/include",
           "-I", os.path.join(ROOT, "libgaussianblur_b200", "csrc"), os.path.join(ROOT, "tests", "cpp", "ols_emulate.cpp"),
           "-o", str(so)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    lib = ctypes.CDLL(str(so))
    lib.ols_emulate_rowpass.restype = ctypes.c_int

    def run(img, sigma, padq=0):
        rows, cols, _ = img.shape
        ksize, pad, n_rp, _, _, _ = go.plan_sizes(rows, cols, sigma)
        img = np.ascontiguousarray(img)
        out = np.empty((rows, cols * 4), np.float32)
        rc = lib.ols_emulate_rowpass(img.ctypes.data_as(ctypes.c_void_p), rows, cols, ctypes.c_float(sigma), ksize, n_rp,
                                     padq, out.ctypes.data_as(ctypes.c_void_p))
        assert rc == 0, f"mirrored border slot mismatch ({rc})"
        return out, ksize, pad, n_rp

    return run


def _oracle_rows(img, sigma, ksize, pad, n):
    want = np.empty((img.shape[0], img.shape[1] * 4), np.float64)
    for ch in range(4):
        want[:, ch::4] = go.conv_lines(img[:, :, ch].astype(np.float64), sigma, ksize, pad, n)
    return want


@pytest.mark.parametrize("rows,cols,sigma,padq", [
    (140, 2000, 20.0, 0),    # three blocks, last one partial
    (150, 3000, 20.0, 18),   # panel slots with bank-padding holes (the 4608-point column plan's padq)
    (40, 700, 5.0, 16),      # one block holds the whole padded row
    (40, 994, 5.0, 0),       # cols == L: exactly one full block
    (40, 995, 5.0, 0),       # one pixel into a second block
    (340, 1500, 50.0, 0),    # pad = 165
    (520, 1100, 76.0, 0),    # pad = 252, close to the widest window the 1024-sample block takes
])
def test_emulated_rowpass_matches_oracle(emu, rows, cols, sigma, padq):
    img = go.synth_u8((rows, cols, 4), 0x51 + cols)
    out, ksize, pad, n = emu(img, sigma, padq)
    want = _oracle_rows(img, sigma, ksize, pad, n)
    assert np.abs(out - want).max() <= TOL


def test_emulated_rowpass_headline_width(emu):
    """the 8K row: 7680 pixels, sigma 20 -> 9 blocks of 894 outputs against the reference's N = 8000"""
    img = go.synth_u8((132, 7680, 4), 0x1236)
    out, ksize, pad, n = emu(img, 20.0, 18)
    assert (ksize, pad) == (131, 65)
    want = _oracle_rows(img, 20.0, ksize, pad, n)
    assert np.abs(out - want).max() <= TOL
    if ref_lib.available():
        ref = ref_lib.rowpass_plane(img, 2, 20.0)
        assert np.abs(out[:, 2::4] - ref).max() <= TOL


@pytest.mark.parametrize("pattern", ["altcols", "altrows", "checker", "const255", "impulse", "ramp"])
def test_emulated_rowpass_adversarial(emu, pattern):
    """alternating columns make the reference's Nyquist term as large as it gets"""
    rows, cols, sigma = 136, 1203, 20.0
    img = np.zeros((rows, cols, 4), np.uint8)
    if pattern == "altcols":
        img[:, ::2] = 255
    elif pattern == "altrows":
        img[::2] = 255
    elif pattern == "checker":
        img[::2, ::2] = 255
        img[1::2, 1::2] = 255
    elif pattern == "const255":
        img[:] = 255
    elif pattern == "impulse":
        img[70, 600] = 255
    else:
        img[:] = (np.arange(cols) % 256).astype(np.uint8)[None, :, None]
    out, ksize, pad, n = emu(img, sigma)
    want = _oracle_rows(img, sigma, ksize, pad, n)
    assert np.abs(out - want).max() <= TOL


@pytest.fixture(scope="module")
def emu_c(tmp_path_factory):
    """ols_row_c1_kernel<.., DIL>: one-channel rows (DIL = 1) and RGB rows run as byte rows (DIL = 3)"""
    so = tmp_path_factory.mktemp("ols3") / "libols_emu3.so"
    cmd = ["g++", "-std=c++17", "-O2", "-shared", "-fPIC", "-w", "-I", "/usr/local/cuda/include",
           "-I", os.path.join(ROOT, "libgaussianblur_b200", "csrc"), os.path.join(ROOT, "tests", "cpp", "ols_emulate.cpp"),
           "-o", str(so)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    lib = ctypes.CDLL(str(so))
    lib.ols_emulate_rowpass_c.restype = ctypes.c_int

    def run(img, sigma, padq=0):
        rows, cols, ch = img.shape
        ksize, pad, n_rp, _, _, _ = go.plan_sizes(rows, cols, sigma)
        img = np.ascontiguousarray(img)
        out = np.empty((rows, cols * ch), np.float32)
        rc = lib.ols_emulate_rowpass_c(img.ctypes.data_as(ctypes.c_void_p), rows, cols, ch, ctypes.c_float(sigma), ksize, n_rp,
                                       padq, out.ctypes.data_as(ctypes.c_void_p))
        assert rc == 0, f"emulation failed ({rc})"
        want = np.empty((rows, cols * ch), np.float64)
        for c in range(ch):
            want[:, c::ch] = go.conv_lines(img[:, :, c].astype(np.float64), sigma, ksize, pad, n_rp)
        return out, want, pad

    return run


@pytest.mark.parametrize("rows,cols,ch,sigma,padq", [
    (70, 1920, 3, 10.0, 0),    # the video workload's row: 5760 bytes, pad 96 bytes, 7 blocks of 832
    (71, 1920, 3, 5.0, 12),    # config 1's row, odd row count (the last pair has one row), slots with holes
    (90, 1280, 3, 20.0, 0),    # pad 65 pixels = 195 bytes
    (80, 4096, 3, 25.0, 0),    # pad 82 pixels = 246 bytes, close to the widest window byte rows take
    (40, 336, 3, 3.0, 0),      # one block holds the whole padded row
    (60, 3840, 3, 1.0, 0),     # pad 2
    (66, 1920, 1, 10.0, 0),    # one channel through the same emulation
    (41, 2016, 1, 20.0, 0),
])
def test_emulated_rowpass_interleaved_and_one_channel(emu_c, rows, cols, ch, sigma, padq):
    img = go.synth_u8((rows, cols, ch), 0x33 + cols)
    out, want, _ = emu_c(img, sigma, padq)
    assert np.abs(out - want).max() <= TOL


@pytest.mark.parametrize("pattern", ["altcols", "altcols_mixed", "const255", "impulse"])
def test_emulated_rowpass_rgb_adversarial(emu_c, pattern):
    """alternating pixel columns, per channel in opposite phase: the per-channel Nyquist terms of the byte-row pass"""
    rows, cols, sigma = 40, 1920, 10.0
    img = np.zeros((rows, cols, 3), np.uint8)
    if pattern == "altcols":
        img[:, ::2] = 255
    elif pattern == "altcols_mixed":
        img[:, ::2, 0] = 255
        img[:, 1::2, 1] = 255
        img[:, ::2, 2] = 128
    elif pattern == "const255":
        img[:] = 255
    else:
        img[20, 960, 1] = 255
    out, want, _ = emu_c(img, sigma)
    assert np.abs(out - want).max() <= TOL


@pytest.fixture(scope="module")
def emu_blur(emu, tmp_path_factory):
    so = tmp_path_factory.mktemp("ols2") / "libols_emu2.so"
    cmd = ["g++", "-std=c++17", "-O2", "-shared", "-fPIC", "-w", "-I", "/usr/local/cuda/include",
           "-I", os.path.join(ROOT, "libgaussianblur_b200", "csrc"), os.path.join(ROOT, "tests", "cpp", "ols_emulate.cpp"),
           "-o", str(so)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    lib = ctypes.CDLL(str(so))
    lib.ols_emulate_blur.restype = ctypes.c_int

    lib.ols_emulate_blur_keep_alpha.restype = ctypes.c_int

    def run(img, sigma, keep_alpha=False):
        rows, cols, _ = img.shape
        ksize, pad, n_rp, n_cp, _, _ = go.plan_sizes(rows, cols, sigma)
        img = np.ascontiguousarray(img)
        out = np.empty_like(img)
        fn = lib.ols_emulate_blur_keep_alpha if keep_alpha else lib.ols_emulate_blur
        rc = fn(img.ctypes.data_as(ctypes.c_void_p), rows, cols, ctypes.c_float(sigma), ksize, n_rp, n_cp,
                out.ctypes.data_as(ctypes.c_void_p))
        assert rc == 0, rc
        return out

    return run


@pytest.mark.parametrize("rows,cols,sigma", [
    (1100, 72, 20.0),     # two column blocks, last one partial
    (2000, 68, 20.0),     # three column blocks
    (300, 132, 5.0),      # one column block
    (1900, 172, 50.0),    # pad = 165
    (895, 68, 20.0),      # one row into a second column block
    (1788, 68, 20.0),     # rows == 2 L: the last block ends exactly at the image
])
def test_emulated_two_pass_blur_matches_oracle(emu_blur, rows, cols, sigma):
    """row pass + column pass (synthetic-row Nyquist term, swizzled TMA landing area, shared u8 tile) against the float64
    oracle of the whole reference call and the compiled reference: max |delta| <= 1 LSB"""
    img = go.synth_u8((rows, cols, 4), 0x77 + rows)
    out = emu_blur(img, sigma)
    want = go.gaussianblur(img, sigma, True)
    assert np.abs(out.astype(int) - want.astype(int)).max() <= 1
    if ref_lib.available():
        ref = ref_lib.gaussianblur(img, sigma, True)
        assert np.abs(out.astype(int) - ref.astype(int)).max() <= 1


@pytest.mark.parametrize("pattern", ["altrows", "altcols", "checker"])
def test_emulated_two_pass_blur_adversarial(emu_blur, pattern):
    """alternating rows make the column Nyquist term as large as it gets"""
    rows, cols, sigma = 1000, 72, 20.0
    img = np.zeros((rows, cols, 4), np.uint8)
    if pattern == "altrows":
        img[::2] = 255
    elif pattern == "altcols":
        img[:, ::2] = 255
    else:
        img[::2, ::2] = 255
        img[1::2, 1::2] = 255
    out = emu_blur(img, sigma)
    want = go.gaussianblur(img, sigma, True)
    assert np.abs(out.astype(int) - want.astype(int)).max() <= 1


@pytest.mark.parametrize("rows,cols,sigma", [(1100, 72, 20.0), (895, 68, 20.0), (300, 132, 5.0)])
def test_emulated_blur_with_alpha_passed_through(emu_blur, rows, cols, sigma):
    """RGBA with apply_to_alpha = false on the block kernels (ols_col_kernel<true, true>): all four channels are
    transformed, the item's strip of the input image lands in the output tile first and the (B, A) warps write only
    their B byte — alpha comes back bit-identical (the reference's own check, tests/test_gaussianblur.cpp:115-118), the
    colour channels within 1 LSB of the reference's three-channel result"""
    img = go.synth_u8((rows, cols, 4), 0x3A + rows)
    out = emu_blur(img, sigma, keep_alpha=True)
    assert np.array_equal(out[..., 3], img[..., 3])
    want = go.gaussianblur(img, sigma, False)
    assert np.abs(out[..., :3].astype(int) - want[..., :3].astype(int)).max() <= 1
    if ref_lib.available():
        ref = ref_lib.gaussianblur(img, sigma, False)
        assert np.array_equal(ref[..., 3], img[..., 3])
        assert np.abs(out.astype(int) - ref.astype(int)).max() <= 1
