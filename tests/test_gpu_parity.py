"""GPU parity tests (pytest -m gpu): the CUDA path, called through the C ABI, against
(1) the golden vectors minted from the compiled reference, (2) the compiled reference itself when
oracle/_ref travelled with the repo (else the pinned numpy restatement), on the same seeded inputs.
Bar (BASELINE.json north_star): max |delta| <= 1 LSB per 8-bit channel; fp32 row-pass
intermediate within 5e-4 absolute; kernel spectra within 1e-4 of max|K|.
Nothing here reads /root/reference."""
import numpy as np
import pytest

from oracle import gblur_oracle as go
from tests.cases import DIGEST_CASES, SMALL_CASES, make_input

pytestmark = pytest.mark.gpu

LSB = 1
INTER_TOL = 5e-4


def _blur_gpu(gb, img, sigma, alpha):
    ch = 1 if img.ndim == 2 else img.shape[2]
    plan = gb.Plan(img.shape[0], img.shape[1], ch, sigma)
    out = img.copy()
    plan.run_host(out, alpha)
    inter = plan.dump_intermediate(alpha)
    rad = (plan.radices(0), plan.radices(1))
    plan.close()
    return out, inter, rad


def _assert_close_u8(out, want, what):
    d = np.abs(out.astype(np.int16) - want.astype(np.int16))
    assert d.max() <= LSB, f"{what}: max |delta| = {d.max()} LSB at {np.unravel_index(d.argmax(), d.shape)}"
    return float((d > 0).mean())


def test_native_library_is_what_runs(gb):
    import torch

    assert torch.cuda.is_available()
    assert gb.lib().gblur_device_count() >= 1
    before = gb.launch_count()
    img = go.synth_u8((32, 32, 3), 1)
    gb.gaussianblur(img, 2.0, False)
    assert gb.launch_count() - before == 2  # row pass + column pass


@pytest.mark.parametrize("case", SMALL_CASES, ids=[c[0] for c in SMALL_CASES])
def test_small_cases_vs_golden(gb, golden_small, case):
    name, rows, cols, ch, sigma, alpha, pat = case
    img = make_input(rows, cols, ch, pat)
    out, inter, _ = _blur_gpu(gb, img, sigma, alpha)
    _assert_close_u8(out, golden_small[name], name)
    cp = 3 if (ch == 4 and not alpha) else ch
    got = inter.reshape(rows, cols, cp)[..., 0]
    assert np.abs(got - golden_small[name + "__rowpass_ch0"]).max() < INTER_TOL
    if ch == 4 and not alpha:
        assert np.array_equal(out[..., 3], img[..., 3])  # alpha untouched bit for bit


def test_gray_plane_vs_golden(gb, golden_small):
    plane = go.synth_u8((48, 64), 0x1234 + 3)
    out, _, _ = _blur_gpu(gb, plane, 4.0, False)
    _assert_close_u8(out, golden_small["gray_48x64_s4"], "gray")


@pytest.mark.parametrize("case", DIGEST_CASES, ids=[c[0] for c in DIGEST_CASES])
def test_medium_cases_vs_reference(gb, refcheck, golden_digests, case):
    name, rows, cols, ch, sigma, alpha, pat = case
    img = make_input(rows, cols, ch, pat)
    out, inter, rad = _blur_gpu(gb, img, sigma, alpha)
    want = refcheck.blur(img, sigma, alpha)
    frac = _assert_close_u8(out, want, name)
    assert frac < 0.02, f"{name}: {frac:.4f} of the bytes differ by 1 LSB (expected only rounding-boundary flips)"
    rec = golden_digests["outputs"][name]
    step = max(1, out.size // 64)
    sample = out.reshape(-1)[::step][:64].astype(int)
    assert np.abs(sample - np.array(rec["sample"])).max() <= LSB
    assert abs(int(out.astype(np.uint64).sum()) - rec["sum"]) <= out.size * 0.02
    cp = 3 if (ch == 4 and not alpha) else ch
    got = inter.reshape(rows, cols, cp)[..., 1]
    assert np.abs(got - refcheck.rowpass(img, 1, sigma)).max() < INTER_TOL


@pytest.mark.parametrize("sigma", [0.3, 1.0, 3.0, 5.0, 20.0, 50.0])
@pytest.mark.parametrize("pat", ["synth", "altcols", "altrows", "checker", "const255", "impulse", "hramp", "vramp"])
def test_adversarial_patterns(gb, refcheck, pat, sigma):
    rows, cols, ch = 270, 480, 4
    img = make_input(rows, cols, ch, pat)
    out, _, _ = _blur_gpu(gb, img, sigma, True)
    _assert_close_u8(out, refcheck.blur(img, sigma, True), f"{pat} sigma={sigma}")


@pytest.mark.parametrize("shape", [(3, 3, 3), (4, 4, 3), (5, 7, 4), (17, 33, 3), (33, 17, 4), (31, 97, 3), (1, 40, 3),
                                   (40, 1, 3), (2, 2, 4), (129, 255, 3), (200, 333, 4)])
def test_small_odd_shapes(gb, refcheck, shape):
    rows, cols, ch = shape
    img = go.synth_u8(shape, 99)
    for sigma in (0.5, 1.0, 2.0):
        sz = gb.host_plan_sizes(rows, cols, sigma)
        if sz["pad"] > min(rows, cols) - 1:
            with pytest.raises(gb.GblurError) as ei:  # reference precondition (it reads out of bounds)
                gb.Plan(rows, cols, ch, sigma)
            assert ei.value.code == 2
            continue
        for alpha in (False, True):
            out, _, _ = _blur_gpu(gb, img, sigma, alpha)
            _assert_close_u8(out, refcheck.blur(img, sigma, alpha), f"{shape} sigma={sigma} alpha={alpha}")


@pytest.mark.parametrize("n_target", [96, 160, 192, 288, 320, 480, 576, 800, 864, 960, 1440, 2400, 2592, 4000, 6912])
def test_many_fft_lengths_runtime_radix_kernel(gb, refcheck, n_target):
    """Line lengths chosen so the row pass lands on FFT length n_target (mixed radices 2/3/5)."""
    sigma = 2.0
    pad = (go.gaussian_window(sigma, 10000) - 1) // 2
    cols = n_target - 2 * pad
    rows = 24
    assert gb.host_plan_sizes(rows, cols, sigma)["n_rowpass"] == n_target
    for ch in (3, 4):
        img = go.synth_u8((rows, cols, ch), n_target)
        out, _, _ = _blur_gpu(gb, img, sigma, True)
        _assert_close_u8(out, refcheck.blur(img, sigma, True), f"N={n_target} ch={ch}")


def test_single_sequence_kernel_long_lines(gb, refcheck):
    """FFT length 17280 does not fit two sequences in shared memory -> one-sequence kernel."""
    sigma = 3.0
    cols = 17280 - 2 * ((go.gaussian_window(sigma, 20000) - 1) // 2)
    lp = gb.host_line_plan(17280)
    assert lp["nseq"] == 1
    img = go.synth_u8((40, cols, 3), 5)
    out, _, _ = _blur_gpu(gb, img, sigma, False)
    _assert_close_u8(out, refcheck.blur(img, sigma, False), "N=17280")


def test_spectra_within_1e4(gb, refcheck):
    from oracle import ref_lib

    for rows, cols, sigma in [(270, 480, 5.0), (1080, 1920, 5.0), (1080, 1920, 10.0)]:
        plan = gb.Plan(rows, cols, 3, sigma)
        mine0, mine1 = plan.spectrum(0), plan.spectrum(1)
        plan.close()
        ks, pad, n_rp, n_cp, *_ = go.plan_sizes(rows, cols, sigma)
        if ref_lib.available():
            k = ref_lib.kernel_dft(rows, cols, sigma)
            want0, want1 = k["kerf_col"], k["kerf_row"]
        else:
            want0, want1 = go.kerf_ordered(sigma, ks, n_rp), go.kerf_ordered(sigma, ks, n_cp)
        assert np.abs(mine0 - want0).max() < 1e-4 and np.abs(mine1 - want1).max() < 1e-4
        big = np.abs(want0) > 1e-3
        assert (np.abs(mine0 - want0)[big] / np.abs(want0)[big]).max() < 1e-4


def test_device_buffers_in_place_and_out_of_place(gb, refcheck):
    import torch

    img = go.synth_u8((270, 480, 4), 3)
    want = refcheck.blur(img, 5.0, False)
    plan = gb.Plan(270, 480, 4, 5.0)
    d_in = torch.from_numpy(img).cuda()
    d_out = torch.zeros_like(d_in)
    plan.run_device(d_in, d_out, apply_to_alpha=False)
    torch.cuda.synchronize()
    assert np.array_equal(d_in.cpu().numpy(), img)  # input untouched
    _assert_close_u8(d_out.cpu().numpy(), want, "out of place")
    assert np.array_equal(d_out.cpu().numpy()[..., 3], img[..., 3])
    plan.run_device(d_in, None, apply_to_alpha=False)  # in place
    torch.cuda.synchronize()
    assert np.array_equal(d_in.cpu().numpy(), d_out.cpu().numpy())
    plan.close()


def test_batch_matches_per_frame(gb, refcheck):
    frames = np.stack([go.synth_u8((135, 240, 3), 0x1234 + 2 + f) for f in range(5)])
    plan = gb.Plan(135, 240, 3, 10.0, max_frames=5)
    out = frames.copy()
    plan.run_host(out, False, n_frames=5)
    plan.close()
    for f in range(5):
        _assert_close_u8(out[f], refcheck.blur(frames[f], 10.0, False), f"frame {f}")


def test_repeat_runs_are_bit_identical(gb):
    img = go.synth_u8((270, 480, 4), 11)
    plan = gb.Plan(270, 480, 4, 5.0)
    a = plan.run_host(img.copy(), True)
    b = plan.run_host(img.copy(), True)
    plan.close()
    assert np.array_equal(a, b)


def test_band_passes_equal_whole_image(gb):
    """Row bands through gblur_rowpass_device + column bands through gblur_colpass_device give
    exactly the bytes of the single-GPU run (the band-sharded multi-GPU path, on one GPU)."""
    import torch

    rows, cols, ch, sigma, g = 96, 160, 3, 4.0, 4
    img = go.synth_u8((rows, cols, ch), 21)
    whole = gb.Plan(rows, cols, ch, sigma).run_host(img.copy(), False)
    w = cols * ch
    inter = torch.zeros((rows, w), dtype=torch.float32, device="cuda")
    d_img = torch.from_numpy(img).cuda()
    br = rows // g
    for k in range(g):
        p = gb.Plan(rows, cols, ch, sigma, band_rows=br, band_cols=cols)
        p.rowpass(d_img[k * br:(k + 1) * br], inter[k * br:(k + 1) * br], w)
        torch.cuda.synchronize()
        p.close()
    out = torch.zeros((rows, cols, ch), dtype=torch.uint8, device="cuda")
    bc = cols // g
    for k in range(g):
        p = gb.Plan(rows, cols, ch, sigma, band_rows=rows, band_cols=bc)
        band_inter = inter[:, k * bc * ch:(k + 1) * bc * ch].contiguous()
        band_out = torch.zeros((rows, bc, ch), dtype=torch.uint8, device="cuda")
        p.colpass(band_inter, bc * ch, band_out, bc * ch)
        torch.cuda.synchronize()
        out[:, k * bc:(k + 1) * bc] = band_out
        p.close()
    assert np.array_equal(out.cpu().numpy(), whole)


@pytest.mark.parametrize("n_target,ch", [(34560, 1), (34560, 3), (69120, 3), (69120, 4)])
def test_cluster_kernel_long_rows(gb, refcheck, n_target, ch):
    """Row FFT lengths of BASELINE.json configs 4 and 5 (34560, 69120) do not fit one CTA: the line is
    split over a thread-block cluster (distributed shared memory)."""
    sigma = 3.0
    pad = (go.gaussian_window(sigma, 100000) - 1) // 2
    cols = n_target - 2 * pad
    rows = 20
    assert gb.host_plan_sizes(rows, cols, sigma)["n_rowpass"] == n_target
    shape = (rows, cols) if ch == 1 else (rows, cols, ch)
    img = go.synth_u8(shape, n_target + ch)
    out, _, _ = _blur_gpu(gb, img, sigma, True)
    _assert_close_u8(out, refcheck.blur(img, sigma, True), f"N={n_target} ch={ch}")


@pytest.mark.parametrize("ch", [1, 3])
def test_cluster_kernel_long_columns(gb, refcheck, ch):
    """Column FFT length 34560 (configs 4 and 5): cluster kernel on the column pass."""
    sigma = 3.0
    pad = (go.gaussian_window(sigma, 100000) - 1) // 2
    rows = 34560 - 2 * pad
    cols = 24
    assert gb.host_plan_sizes(rows, cols, sigma)["n_colpass"] == 34560
    shape = (rows, cols) if ch == 1 else (rows, cols, ch)
    img = go.synth_u8(shape, 77 + ch)
    out, _, _ = _blur_gpu(gb, img, sigma, True)
    _assert_close_u8(out, refcheck.blur(img, sigma, True), f"N_col=34560 ch={ch}")


def test_large_sigma_long_line(gb, refcheck):
    """sigma=50 on a 33000-wide single plane (config 4's kernel: 331 taps, N=34560 on rows)."""
    sigma = 50.0
    sz = gb.host_plan_sizes(400, 34000, sigma)
    assert sz["n_rowpass"] == 34560
    img = go.synth_u8((400, 34000), 4)
    out, _, _ = _blur_gpu(gb, img, sigma, False)
    _assert_close_u8(out, refcheck.blur(img, sigma, False), "sigma=50 N=34560")


def test_band_sharded_binding_single_rank(gb, refcheck):
    """libgaussianblur_b200.multi.BandShardedBlur bound to the CUDA band plans, world size 1."""
    import torch

    from libgaussianblur_b200.multi import BandShardedBlur

    for ch, alpha, fused in ((1, False, True), (3, False, True), (4, False, True), (4, True, True), (3, False, False),
                             (4, False, False)):
        rows, cols, sigma = 96, 160, 4.0
        img = go.synth_u8((rows, cols, ch) if ch > 1 else (rows, cols), 31 + ch)
        b = BandShardedBlur.cuda(rows, cols, ch, sigma, alpha, fused=fused)
        assert (b.fused_run is not None) == fused
        out = b.run(torch.from_numpy(img).cuda().view(rows, cols, ch))
        torch.cuda.synchronize()
        b.close()
        want = refcheck.blur(img, sigma, alpha)
        _assert_close_u8(out.cpu().numpy().reshape(want.shape), want, f"band binding ch={ch} alpha={alpha}")


def test_sigma150_rows_69120(gb, refcheck):
    """Config 5's row pass: sigma=150 (997 taps, pad 498) on 68124-wide RGB rows -> FFT length 69120."""
    sigma = 150.0
    rows, cols = 600, 69120 - 2 * 498
    sz = gb.host_plan_sizes(rows, cols, sigma)
    assert sz["pad"] == 498 and sz["n_rowpass"] == 69120
    img = go.synth_u8((rows, cols, 3), 150)
    out, _, _ = _blur_gpu(gb, img, sigma, False)
    _assert_close_u8(out, refcheck.blur(img, sigma, False), "sigma=150 N=69120")


@pytest.mark.parametrize("ch,alpha", [(4, True), (4, False), (3, False)])
def test_tma_column_gather_n4608(gb, refcheck, ch, alpha):
    """Column FFT length 4608 (the 8K headline shape): compile-time kernel whose panel input comes through
    strided TMA tiles, with the next unit requested during the last inverse stage."""
    rows, cols, sigma = 4320, 200, 20.0
    sz = gb.host_plan_sizes(rows, cols, sigma)
    assert sz["n_colpass"] == 4608 and gb.host_line_plan(4608)["static"]
    img = go.synth_u8((rows, cols, ch), 4608 + ch)
    out, _, _ = _blur_gpu(gb, img, sigma, alpha)
    _assert_close_u8(out, refcheck.blur(img, sigma, alpha), f"N_col=4608 ch={ch} alpha={alpha}")
    if ch == 4 and not alpha:
        assert np.array_equal(out[..., 3], img[..., 3])


@pytest.mark.parametrize("ch,alpha", [(4, True), (3, False), (4, False)])
def test_pipelined_host_path_equals_device_path(gb, ch, alpha):
    """gblur_run_host on a large image goes through the PCIe-pipelined path (row chunks under the H2D copy,
    column bands under the D2H copy): the result must equal the plain device-buffer run."""
    import torch

    rows, cols, sigma = 1500, 2500, 7.0
    img = go.synth_u8((rows, cols, ch), 900 + ch)
    assert img.nbytes >= (8 << 20)
    plan = gb.Plan(rows, cols, ch, sigma)
    d_in = torch.from_numpy(img).cuda()
    d_out = torch.empty_like(d_in)
    plan.run_device(d_in, d_out, alpha)
    torch.cuda.synchronize()
    want = d_out.cpu().numpy()
    pinned = torch.from_numpy(img.copy()).pin_memory()
    plan.run_host(pinned, alpha)
    got_pinned = pinned.numpy().copy()
    pageable = plan.run_host(img.copy(), alpha)
    plan.close()
    assert np.array_equal(got_pinned, pageable)
    if ch == 4 and alpha:
        assert np.array_equal(got_pinned, want)  # RGBA: a pixel's own channels share one complex transform
    else:
        if ch == 4:
            assert np.array_equal(got_pinned[..., 3], img[..., 3])  # alpha untouched
        # three processed channels: lines are paired (a + ib) across row boundaries; chunking changes the pairing, hence the last
        # bits of the fp32 intermediate: rare 1-LSB flips at rounding boundaries, nothing else
        d = np.abs(got_pinned.astype(np.int16) - want.astype(np.int16))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3


def test_batch_longer_than_plan_is_streamed(gb, refcheck):
    frames = np.stack([go.synth_u8((135, 240, 3), 0x2234 + f) for f in range(7)])
    plan = gb.Plan(135, 240, 3, 6.0, max_frames=2)
    out = frames.copy()
    plan.run_host(out, False, n_frames=7)
    plan.close()
    for f in range(7):
        _assert_close_u8(out[f], refcheck.blur(frames[f], 6.0, False), f"streamed frame {f}")


@pytest.mark.parametrize("rows,cols,ch,alpha", [(1081, 1920, 3, False), (1082, 1920, 3, False), (1083, 1918, 3, False),
                                                (1080, 1920, 4, False), (271, 1900, 4, False)])
def test_three_channel_row_problem_n2048(gb, refcheck, rows, cols, ch, alpha):
    """Row FFT length 2048 (the 1080p shape) with three processed channels: the specialised row problem
    (two image rows per unit, staged through shared memory when the rows are 16-byte aligned), every tail
    phase of rows*3 mod 4, unaligned rows (no staging), and RGBA with the alpha plane left alone."""
    import os
    sigma = 5.0
    assert gb.host_plan_sizes(rows, cols, sigma)["n_rowpass"] == 2048
    img = go.synth_u8((rows, cols, ch), rows + cols)
    out, _, rad = _blur_gpu(gb, img, sigma, alpha)
    assert rad[0][1], "compile-time row kernel expected"
    _assert_close_u8(out, refcheck.blur(img, sigma, alpha), f"{rows}x{cols}x{ch}")
    if ch == 4:
        assert np.array_equal(out[..., 3], img[..., 3])
    os.environ["GBLUR_NO_RGB3"] = "1"  # same lines paired the same way: the general row problem gives the same bits
    try:
        out2, _, _ = _blur_gpu(gb, img, sigma, alpha)
    finally:
        del os.environ["GBLUR_NO_RGB3"]
    assert np.array_equal(out, out2)


def test_three_channel_row_problem_batch_straddles_frames(gb, refcheck):
    """Frames stacked as extra rows: with 5 rows x 3 channels per frame a unit of four lines crosses frame
    boundaries; runtime-radix kernel (N = 160) and compile-time kernel (N = 2048)."""
    for rows, cols, sigma in ((5, 150, 1.0), (33, 1920, 5.0)):
        frames = 3
        batch = go.synth_u8((frames, rows, cols, 3), 4242 + cols)
        plan = gb.Plan(rows, cols, 3, sigma, max_frames=frames)
        out = batch.copy()
        plan.run_host(out, False, n_frames=frames)
        plan.close()
        for f in range(frames):
            _assert_close_u8(out[f], refcheck.blur(batch[f], sigma, False), f"frame {f} of {rows}x{cols}")


def test_python_zero_copy_device_arrays(gb, refcheck):
    """gaussianblur() on GPU-resident pixels — a torch CUDA tensor and a bare
    __cuda_array_interface__ exporter — blurs the caller's device memory in place, no host round trip."""
    import torch

    img = go.synth_u8((211, 317, 4), 77)
    want = refcheck.blur(img, 4.0, True)
    t = torch.from_numpy(img).cuda()
    assert gb.gaussianblur(t, 4.0, True) is t
    torch.cuda.synchronize()
    _assert_close_u8(t.cpu().numpy(), want, "torch CUDA tensor")

    class Exporter:  # what CuPy / Numba arrays look like to a consumer
        def __init__(self, tensor):
            self._keep = tensor
            self.__cuda_array_interface__ = {"shape": tuple(tensor.shape), "typestr": "|u1", "version": 3,
                                             "data": (tensor.data_ptr(), False), "strides": None}

    t2 = torch.from_numpy(img).cuda()
    torch.cuda.synchronize()
    gb.gaussianblur(Exporter(t2), 4.0, True)
    torch.cuda.synchronize()
    assert torch.equal(t, t2)
    with pytest.raises(TypeError):
        gb.gaussianblur(torch.zeros((8, 8, 3), dtype=torch.float32, device="cuda"), 1.0)
    # reference no-op cases keep their behaviour on device data too
    t3 = torch.from_numpy(img).cuda()
    gb.gaussianblur(t3, 0.0, True)
    assert torch.equal(t3.cpu(), torch.from_numpy(img))
