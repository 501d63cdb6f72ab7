"""GPU parity of the overlap-save row pass on RGB images (csrc/ols_kernels.cu: ols_row_c1_kernel<.., DIL = 3>): a row of
three interleaved channels runs as ONE byte row convolved with the taps spread three bytes apart, so a transform serves
all channels and every store is a contiguous run of floats (csrc/ols_core.cuh: reflect_src_dil, store_block_c1<3>). The
column pass stays the whole-line kernel. Checked through the C ABI against the compiled reference
(src/gaussianblur.cpp:202-244 per channel): max |delta| <= 1 LSB per 8-bit channel, fp32 row-pass intermediate within
5e-4 absolute; `last_kernels` and the launch counts prove which kernels ran (overlap-save row pass = 2 launches:
alternating sums + blocks)."""
import os

import numpy as np
import pytest

from oracle import gblur_oracle as go

pytestmark = pytest.mark.gpu

INTER_TOL = 5e-4


def _plan(gb, rows, cols, sigma, env, max_frames=1):
    saved = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        return gb.Plan(rows, cols, 3, sigma, max_frames=max_frames)
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def _run(gb, img, sigma, env):
    import torch

    plan = _plan(gb, img.shape[0], img.shape[1], sigma, env)
    d_in = torch.from_numpy(img).cuda()
    d_out = torch.empty_like(d_in)
    plan.run_device(d_in, d_out, False)
    torch.cuda.synchronize()
    before = gb.launch_count()
    plan.run_device(d_in, d_out, False)
    torch.cuda.synchronize()
    launches = gb.launch_count() - before
    kinds = plan.last_kernels()
    inter = plan.dump_intermediate(False)
    plan.close()
    return d_out.cpu().numpy(), inter, launches, kinds


def _check(out, inter, img, sigma, refcheck, what):
    want = refcheck.blur(img, sigma, False)
    d = np.abs(out.astype(np.int16) - want.astype(np.int16))
    assert d.max() <= 1, f"{what}: max |delta| = {d.max()} LSB"
    assert (d > 0).mean() < 0.02
    rows, cols, _ = img.shape
    for ch in (0, 1, 2):
        got = inter.reshape(rows, cols, 3)[..., ch]
        assert np.abs(got - refcheck.rowpass(img, ch, sigma)).max() < INTER_TOL, f"{what}: channel {ch}"


@pytest.mark.parametrize("sigma", [10.0, 5.0])
def test_1080p_rgb_overlap_save_row_pass_forced_vs_automatic(gb, refcheck, sigma):
    """BASELINE.json configs 1 and 3: 1920 x 1080 RGB, sigma 5 / 10 — byte rows of 5760, 7 blocks per row pair. The
    automatic rule leaves RGB to the whole-line kernels (measured faster at this size, ols_kernels.cu row_profitable)."""
    img = go.synth_u8((1080, 1920, 3), 0x1234 + int(sigma))
    out, inter, launches, kinds = _run(gb, img, sigma, {"GBLUR_ROW_OLS": "1"})
    assert kinds[0] == "overlap_save_1024" and kinds[1] != "overlap_save_1024", kinds
    assert launches == 3, "alternating sums + row blocks + whole-line column pass expected"
    _check(out, inter, img, sigma, refcheck, f"1080p sigma={sigma}, forced")
    out0, inter0, launches0, kinds0 = _run(gb, img, sigma, {})
    assert launches0 == 2 and kinds0[0] != "overlap_save_1024"
    assert np.abs(out.astype(np.int16) - out0.astype(np.int16)).max() <= 1
    assert np.abs(inter - inter0).max() < INTER_TOL


@pytest.mark.parametrize("rows,cols,sigma", [
    (301, 1360, 3.0),     # odd row count: the last row pair has one row
    (77, 2000, 20.0),     # pad 65 pixels = 195 bytes, fewer rows than two pads
    (130, 4096, 25.0),    # pad 82 pixels = 246 bytes: the widest window byte rows take
    (64, 336, 1.0),       # one block per row
])
def test_forced_overlap_save_rgb_row_pass(gb, refcheck, rows, cols, sigma):
    img = go.synth_u8((rows, cols, 3), rows + cols)
    out, inter, launches, kinds = _run(gb, img, sigma, {"GBLUR_ROW_OLS": "1"})
    assert kinds[0] == "overlap_save_1024", kinds
    _check(out, inter, img, sigma, refcheck, f"{rows}x{cols} sigma={sigma} forced")


def test_rgb_shapes_the_byte_rows_do_not_take_fall_back(gb, refcheck):
    """rows that are not whole 16-byte groups, and windows wider than 256 bytes, keep the whole-line kernels"""
    for rows, cols, sigma in ((90, 1001, 4.0), (400, 1920, 30.0)):
        img = go.synth_u8((rows, cols, 3), 5 + cols)
        out, inter, launches, kinds = _run(gb, img, sigma, {"GBLUR_ROW_OLS": "1"})
        assert kinds[0] != "overlap_save_1024", kinds
        _check(out, inter, img, sigma, refcheck, f"{rows}x{cols} sigma={sigma} fallback")


@pytest.mark.parametrize("pattern", ["altcols_mixed", "checker", "const255"])
def test_rgb_overlap_save_adversarial(gb, refcheck, pattern):
    rows, cols, sigma = 200, 1920, 10.0
    img = np.zeros((rows, cols, 3), np.uint8)
    if pattern == "altcols_mixed":  # per-channel Nyquist terms of opposite sign
        img[:, ::2, 0] = 255
        img[:, 1::2, 1] = 255
        img[:, ::2, 2] = 128
    elif pattern == "checker":
        img[::2, ::2] = 255
        img[1::2, 1::2] = 255
    else:
        img[:] = 255
    out, inter, _, kinds = _run(gb, img, sigma, {"GBLUR_ROW_OLS": "1"})
    assert kinds[0] == "overlap_save_1024"
    _check(out, inter, img, sigma, refcheck, pattern)


def test_rgb_batch_and_host_paths_run_the_overlap_save_row_pass(gb, refcheck):
    """a batch of frames (the video workload's launch shape, odd row count so that row pairs straddle frames) and the
    PCIe-pipelined host call (row chunks) against the per-frame reference"""
    import torch

    rows, cols, sigma, frames = 271, 1920, 10.0, 5
    batch = np.stack([go.synth_u8((rows, cols, 3), 0x900 + f) for f in range(frames)])
    want = np.stack([refcheck.blur(batch[f], sigma, False) for f in range(frames)])
    plan = _plan(gb, rows, cols, sigma, {"GBLUR_ROW_OLS": "1"}, max_frames=frames)
    d_in = torch.from_numpy(batch).cuda()
    d_out = torch.empty_like(d_in)
    plan.run_device(d_in, d_out, False, n_frames=frames)
    torch.cuda.synchronize()
    kinds = plan.last_kernels()
    got = d_out.cpu().numpy()
    plan.close()
    assert kinds[0] == "overlap_save_1024", kinds
    assert np.abs(got.astype(np.int16) - want.astype(np.int16)).max() <= 1
    # host path: one 1080p image (8 row chunks under the H2D copy)
    img = go.synth_u8((1080, 1920, 3), 0x77)
    plan = _plan(gb, 1080, 1920, sigma, {"GBLUR_ROW_OLS": "1"})
    out = plan.run_host(img.copy(), False)
    kinds = plan.last_kernels()
    plan.close()
    assert kinds[0] == "overlap_save_1024", kinds
    assert np.abs(out.astype(np.int16) - refcheck.blur(img, sigma, False).astype(np.int16)).max() <= 1
