"""Parity at the sizes that are actually benchmarked ("Next round" 1a): the CUDA path through the C ABI
against the compiled, unmodified reference (oracle/_ref, conftest.refcheck) on the SAME full-size inputs — the launch
geometry of the headline number (960 panels, 64-bit offsets into the 0.6 GB intermediate, 9 row blocks x 5 column
blocks, TMA tile loads/stores) is the one under test here, not a slice of it.

  * 4320 x 7680 x 4, sigma = 20 (BASELINE.json config 2): alpha blurred and alpha passed through, device-resident
    (gblur_run_device: overlap-save passes) and through the host-buffer call (gblur_run_host: PCIe-pipelined chunks).
  * 10240 x 10240 x 4, sigma = 5: the reference's own stress geometry (/root/reference/tests/test_gaussianblur.cpp:124-164,
    python/tests/stress.py:23-47), which the reference only checks for "variance decreased, < 10 s".
  * one channel, sigma = 50, FFT length 34560 (config 4's cluster kernels) on the row pass (4096 x 34000) and on the
    column pass (34000 x 4096), and 16384 x 16384 (N = 17280) — against ref_lib.blur_plane_u8 = two
    process_channel_tiles calls of the reference (src/gaussianblur.cpp:202-244).
Bar: max |delta| <= 1 LSB per 8-bit channel (BASELINE.json north_star). Reference time on 16 host cores: 1-6 s per case.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cmp(out, want):
    d = np.abs(out.astype(np.int16) - want.astype(np.int16))
    return int(d.max()), int(np.count_nonzero(d))


@pytest.fixture(scope="module")
def headline(refcheck):
    """input and the reference's outputs for both alpha modes (computed once: ~1.5 s each on 16 cores)"""
    from libgaussianblur_b200.synth import synth_u8

    if not refcheck.have_ref:
        pytest.skip("oracle/_ref (compiled reference) did not travel with the repo")
    img = synth_u8((4320, 7680, 4), 0x1234 + 1)
    return img, {True: refcheck.blur(img, 20.0, True), False: refcheck.blur(img, 20.0, False)}


@pytest.mark.parametrize("alpha", [True, False])
def test_headline_8k_device_path_vs_reference(gb, headline, alpha):
    import torch

    img, want = headline
    plan = gb.Plan(4320, 7680, 4, 20.0)
    d_in = torch.from_numpy(img).cuda()
    d_out = torch.empty_like(d_in)
    plan.run_device(d_in, d_out, alpha)
    torch.cuda.synchronize()
    kinds = plan.last_kernels()
    out = d_out.cpu().numpy()
    # in place as well (the column terms of the overlap-save pass read the input the output overwrites)
    plan.run_device(d_in, d_in, alpha)
    torch.cuda.synchronize()
    out_inplace = d_in.cpu().numpy()
    plan.close()
    mx, nd = _cmp(out, want[alpha])
    assert mx <= 1, f"max |delta| = {mx} LSB ({nd} bytes differ), kernels {kinds}"
    assert np.array_equal(out, out_inplace)
    assert kinds == ("overlap_save_1024", "overlap_save_1024"), kinds  # the kernels of the headline number, both modes
    if not alpha:
        assert np.array_equal(out[..., 3], img[..., 3])


@pytest.mark.parametrize("alpha", [True, False])
def test_headline_8k_host_path_vs_reference(gb, headline, alpha):
    img, want = headline
    plan = gb.Plan(4320, 7680, 4, 20.0)
    out = plan.run_host(img.copy(), alpha)
    plan.close()
    mx, nd = _cmp(out, want[alpha])
    assert mx <= 1, f"max |delta| = {mx} LSB ({nd} bytes differ)"
    if not alpha:
        assert np.array_equal(out[..., 3], img[..., 3])


def test_reference_stress_geometry_10240_rgba_vs_reference(gb, refcheck):
    from libgaussianblur_b200.synth import synth_u8

    if not refcheck.have_ref:
        pytest.skip("oracle/_ref did not travel with the repo")
    img = synth_u8((10240, 10240, 4), 0xBEEF)
    want = refcheck.blur(img, 5.0, True)
    plan = gb.Plan(10240, 10240, 4, 5.0)
    out = plan.run_host(img.copy(), True)
    plan.close()
    mx, nd = _cmp(out, want)
    assert mx <= 1, f"max |delta| = {mx} LSB ({nd} bytes differ)"


@pytest.mark.parametrize("rows,cols", [(4096, 34000), (34000, 4096), (16384, 16384)])
def test_single_channel_sigma50_long_lines_vs_reference(gb, refcheck, rows, cols):
    import torch

    from libgaussianblur_b200.synth import synth_u8

    if not refcheck.have_ref:
        pytest.skip("oracle/_ref did not travel with the repo")
    img = synth_u8((rows, cols), 0x1234 + 3)
    want = refcheck.blur(img, 50.0, False)
    plan = gb.Plan(rows, cols, 1, 50.0)
    sizes = (plan.n_rowpass, plan.n_colpass)
    d_in = torch.from_numpy(img).cuda().reshape(rows, cols, 1)
    d_out = torch.empty_like(d_in)
    plan.run_device(d_in, d_out, False)
    torch.cuda.synchronize()
    out = d_out.cpu().numpy().reshape(rows, cols)
    plan.close()
    if rows == 16384:
        assert sizes == (17280, 17280)
    else:
        assert 34560 in sizes
    mx, nd = _cmp(out, want)
    assert mx <= 1, f"max |delta| = {mx} LSB ({nd} bytes differ), FFT lengths {sizes}"


def test_config5_own_geometry_strips_vs_reference(gb, refcheck):
    """BASELINE.json config 5 at its own size: 32768 rows x 65536 columns x RGB, sigma = 150 (pad 498, row FFT 69120,
    column FFT 34560), outside the reference's int rows*cols envelope. The oracle is oracle/ref_harness.cpp's
    ref_band_strip_u8: pass 1 in row bands, pass 2 on a strip of columns, through the reference's own
    process_channel_tiles (src/gaussianblur.cpp:202-244, driven as :261-280) — bit-identical to the monolithic path where
    that one exists (tests/test_oracle.py). Two 64-column strips of one channel are compared (left border and interior):
    every pixel of a strip went through a full-length row transform and a full-length column transform."""
    import torch

    from oracle import ref_lib

    if not refcheck.have_ref:
        pytest.skip("oracle/_ref did not travel with the repo")
    free, _ = torch.cuda.mem_get_info()
    if free < 60e9:
        pytest.skip("needs ~45 GB of device memory")
    rows, cols, sigma = 32768, 65536, 150.0
    rng = np.random.default_rng(0x1234 + 5)
    plane = rng.integers(0, 256, size=(rows, cols), dtype=np.uint8)
    d_plane = torch.from_numpy(plane).cuda()
    d_in = torch.empty((rows, cols, 3), dtype=torch.uint8, device="cuda")
    d_in[..., 0] = 255 - d_plane          # the other two channels: different content, same statistics
    d_in[..., 1] = d_plane                # the channel that is checked
    d_in[..., 2] = torch.roll(d_plane, 7, 1)
    del d_plane
    plan = gb.Plan(rows, cols, 3, sigma)
    assert (plan.pad, plan.n_rowpass, plan.n_colpass) == (498, 69120, 34560)
    d_out = torch.empty_like(d_in)
    plan.run_device(d_in, d_out, False)
    torch.cuda.synchronize()
    worst = 0
    for c0 in (0, 40000):
        got = d_out[:, c0:c0 + 64, 1].cpu().numpy()
        want = ref_lib.band_strip_u8(plane, sigma, c0, 64, band_rows=1024)
        mx, nd = _cmp(got, want)
        worst = max(worst, mx)
        assert mx <= 1, f"strip at column {c0}: max |delta| = {mx} LSB ({nd} bytes differ)"
    plan.close()
    assert worst <= 1
