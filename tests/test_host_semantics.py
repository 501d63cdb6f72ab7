"""CPU tests of the host layer behind the C ABI: the library loads without a GPU, exports every
symbol include/gblur_capi.h declares, reproduces the reference's window / FFT-length / tap /
spectrum semantics, and plans stages whose algebra (checked by a numpy emulation of the kernel's
decimation-in-frequency / decimation-in-time flow) reproduces the oracle. No compute call is made."""
import ctypes
import os
import re

import numpy as np
import pytest

from oracle import gblur_oracle as go
from tests.cases import PLAN_CASES

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(gb):
    hdr = open(os.path.join(ROOT, "include", "gblur_capi.h")).read()
    declared = set(re.findall(r"\b(gblur_[a-z_0-9]+)\s*\(", hdr))
    declared -= {"gblur_plan"}
    assert declared == set(gb.EXPORTED_SYMBOLS), declared ^ set(gb.EXPORTED_SYMBOLS)
    L = ctypes.CDLL(gb.build.lib_path())
    for name in declared:
        assert hasattr(L, name), name


@pytest.mark.parametrize("rows,cols,sigma", PLAN_CASES)
def test_plan_sizes(gb, rows, cols, sigma, golden_digests):
    want = golden_digests["plans"][f"{rows}x{cols}_s{sigma}"]
    got = gb.host_plan_sizes(rows, cols, sigma)
    assert {k: got[k] for k in want} == want
    assert got["ksize"] == go.gaussian_window(sigma, max(rows, cols))


@pytest.mark.parametrize("sigma", [0.3, 1.0, 3.0, 5.0, 20.0, 50.0, 150.0])
def test_taps_bit_identical_to_restatement(gb, sigma):
    ks = go.gaussian_window(sigma)
    assert np.array_equal(gb.host_taps(sigma, ks), go.gaussian_taps(sigma, ks))


def test_spectrum_against_reference_samples(gb, golden_digests):
    for key, rec in golden_digests["spectra"].items():
        dims, s = key.split("_s")
        rows, cols = map(int, dims.split("x"))
        sz = gb.host_plan_sizes(rows, cols, float(s))
        mine = gb.host_spectrum(float(s), sz["ksize"], sz["n_rowpass"])
        for i, v in zip(rec["idx"], rec["kerf_col"]):
            assert abs(float(mine[i]) - v) < 1e-4


def _valid_sizes(limit):
    n, out = 32, []
    while n <= limit:
        if go.is_valid_size(n):
            out.append(n)
        n += 32
    return out


def test_every_valid_length_gets_a_plan(gb):
    for n in _valid_sizes(27000):
        lp = gb.host_line_plan(n)
        assert int(np.prod(lp["radices"])) == n
        assert 2 <= len(lp["radices"]) <= 5
        assert lp["smem_bytes"] <= 227 * 1024
        assert lp["nseq"] in (1, 2)
        if not lp["static"]:
            assert all(r in (2, 3, 4, 5, 6, 8, 9, 10, 12, 15, 16) for r in lp["radices"])


def test_headline_lengths_use_compile_time_kernels(gb):
    for n, rad in [(8000, [20, 20, 20]), (4608, [16, 16, 18]), (2048, [8, 16, 16]), (1152, [12, 12, 8])]:
        lp = gb.host_line_plan(n)
        assert lp["static"] and lp["radices"] == rad and lp["nseq"] == 2


def _emulate_line(x, radices, kperm):
    """numpy model of run_line_*: in-place DIF stages with post-twiddles, gains in position order,
    in-place DIT stages with conjugate pre-twiddles (libgaussianblur_b200/csrc/fftconv.cuh)."""
    n = len(x)
    z = x.astype(np.complex128).copy()
    ms, m = [], n
    for r in radices:
        m //= r
        ms.append(m)

    def stage(r, m, inv, last):
        j = np.arange(r)
        f = np.exp((2j if inv else -2j) * np.pi * np.outer(j, j) / r)
        for u in range(n // r):
            b, mm = divmod(u, m)
            idx = b * r * m + mm + j * m
            c = z[idx]
            tw = np.exp(-2j * np.pi * mm * j / (r * m))
            if inv and not last:
                c = c * np.conj(tw)
            c = f @ c
            if not inv and not last:
                c = c * tw
            z[idx] = c

    for s in range(len(radices) - 1):
        stage(radices[s], ms[s], False, False)
    stage(radices[-1], 1, False, True)
    z *= kperm
    stage(radices[-1], 1, True, True)
    for s in range(len(radices) - 2, -1, -1):
        stage(radices[s], ms[s], True, False)
    return z


@pytest.mark.parametrize("t,sigma", [(33, 1.0), (100, 3.0), (480, 5.0), (1000, 12.0)])
def test_stage_algebra_and_permuted_spectrum(gb, t, sigma):
    ks, pad, n, *_ = go.plan_sizes(t, t, sigma)
    lp = gb.host_line_plan(n)
    kp = gb.host_kperm(sigma, ks, n).astype(np.float64)
    rng = np.random.default_rng(t)
    a = rng.integers(0, 256, t).astype(np.float64)
    b = np.where(np.arange(t) % 2 == 0, 0.0, 255.0)  # alternating line exercises the Nyquist bin
    z = _emulate_line(go.pad_reflect101(a, pad, n) + 1j * go.pad_reflect101(b, pad, n), lp["radices"], kp)
    ra = go.conv_lines(a[None, :], sigma, ks, pad, n)[0]
    rb = go.conv_lines(b[None, :], sigma, ks, pad, n)[0]
    assert np.abs(z.real[pad:pad + t] - ra).max() < 1e-4
    assert np.abs(z.imag[pad:pad + t] - rb).max() < 1e-4


def test_no_gpu_means_loud_failure_not_fallback(gb):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(gb.GblurError) as ei:
        gb.Plan(64, 64, 3, 2.0)
    assert ei.value.code == 4  # GBLUR_E_CUDA
    img = go.synth_u8((16, 16, 3), 7)
    with pytest.raises(gb.GblurError):
        gb.gaussianblur(img.copy(), 2.0, False)


def test_pybind_module_surface(gb):
    m = gb.gaussianblur_module()
    g = m.ImgGeom()
    g.rows, g.cols, g.channels = 2, 2, 2
    img = m.Image()
    img.geom = g
    img.data = [1, 2, 3, 4, 5, 6, 7, 8]
    # unsupported channel count and sigma <= 0: message, data untouched, no GPU needed
    m.gaussianblur(img, 2.0, False)
    assert list(img.data) == [1, 2, 3, 4, 5, 6, 7, 8]
    g.channels = 3
    img.geom = g
    img.data = list(range(12))
    m.gaussianblur(image=img, sigma=0.0, apply_to_alpha=False)
    assert list(img.data) == list(range(12))


def test_automatic_kernel_choice_table():
    """gblur_host_kernel_choice: the rule that picks overlap-save blocks (2 = both passes, 1 = row pass only) or
    whole-line transforms (0) for a single-frame run, pinned at the shapes it was measured on (B200,
    profiles/r02/gates.jsonl, auto_rule.jsonl, gates_c1.jsonl): GPU runs report the same through last_kernels()."""
    import libgaussianblur_b200 as gb

    table = [
        ((4320, 7680, 4, 20.0, True), 2),     # the headline image
        ((4320, 7680, 4, 20.0, False), 2),    # alpha passed through: four channels through the blocks + alpha restore
        ((2160, 3840, 4, 10.0, True), 2),     # 0.370 -> 0.181 ms
        ((2160, 3840, 4, 20.0, True), 2),     # blocks transform 1.28 x / 1.33 x the reference's points: still 2 x faster
        ((1440, 2560, 4, 10.0, True), 2),     # 0.179 -> 0.095 ms
        ((1080, 1920, 4, 5.0, True), 0),      # below 3 Mpixel: whole lines (0.0615 ms against 0.0712)
        ((1080, 1920, 3, 10.0, False), 0),    # RGB: byte rows are forced-only
        ((160, 7680, 4, 20.0, True), 1),      # too few rows for column blocks: wide rows still run as blocks
        ((160, 7680, 4, 20.0, False), 0),
        ((32768, 32768, 1, 50.0, False), 2),  # BASELINE.json config 4
        ((4096, 4096, 1, 50.0, False), 2),    # 0.193 -> 0.127 ms
        ((32768, 65536, 3, 150.0, False), 0), # config 5: pad 498 does not fit a 1024-sample block
        ((8192, 8192, 4, 150.0, True), 0),
        ((4320, 7681, 4, 20.0, True), 1),     # rows of 30724 bytes, not whole 16-byte groups: no TMA column blocks
        ((4320, 7681, 1, 20.0, False), 0),
    ]
    for args, want in table:
        assert gb.host_kernel_choice(*args) == want, args
