"""GPU parity of the overlap-save kernels (csrc/ols_kernels.cu) through the C ABI, against the compiled reference
(or the pinned oracle): max |delta| <= 1 LSB per 8-bit channel, fp32 row-pass intermediate within 5e-4 absolute.
The row pass is what the library picks by itself for wide RGBA rows (7680 pixels, sigma 20: the headline shape);
GBLUR_ROW_OLS=1 / GBLUR_COL_OLS=1 (read when the plan is created) force the kernels on shapes the automatic rule
leaves to the whole-line kernels. The launch counts prove which kernels ran: whole-line = 1 per pass, overlap-save
row pass = 2 (alternating sums + blocks), overlap-save column pass = 3 (column sums, synthetic row, blocks).
Runs last (file name): these kernels are new in round 1."""
import os

import numpy as np
import pytest

from oracle import gblur_oracle as go

pytestmark = pytest.mark.gpu

INTER_TOL = 5e-4


def _blur_device(gb, img, sigma, env):
    import torch

    saved = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        plan = gb.Plan(img.shape[0], img.shape[1], 4, sigma)
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    d_in = torch.from_numpy(img).cuda()
    d_out = torch.empty_like(d_in)
    plan.run_device(d_in, d_out, True)  # warm-up: workspaces, lazily allocated sums
    torch.cuda.synchronize()
    before = gb.launch_count()
    plan.run_device(d_in, d_out, True)
    torch.cuda.synchronize()
    launches = gb.launch_count() - before
    inter = plan.dump_intermediate(True)
    plan.close()
    return d_out.cpu().numpy(), inter, launches


def _check(out, inter, img, sigma, refcheck, what):
    want = refcheck.blur(img, sigma, True)
    d = np.abs(out.astype(np.int16) - want.astype(np.int16))
    assert d.max() <= 1, f"{what}: max |delta| = {d.max()} LSB"
    assert (d > 0).mean() < 0.02
    rows, cols, _ = img.shape
    got = inter.reshape(rows, cols, 4)[..., 2]
    assert np.abs(got - refcheck.rowpass(img, 2, sigma)).max() < INTER_TOL, what


def test_headline_width_picks_overlap_save_row_pass(gb, refcheck):
    img = go.synth_u8((200, 7680, 4), 0x0150)
    out, inter, launches = _blur_device(gb, img, 20.0, {})
    assert launches == 3, "alternating sums + row blocks + whole-line column pass expected"
    _check(out, inter, img, 20.0, refcheck, "7680 wide, automatic")
    out0, inter0, launches0 = _blur_device(gb, img, 20.0, {"GBLUR_ROW_OLS": "0"})
    assert launches0 == 2
    assert np.abs(out.astype(np.int16) - out0.astype(np.int16)).max() <= 1
    assert np.abs(inter - inter0).max() < INTER_TOL


@pytest.mark.parametrize("rows,cols,sigma", [(300, 4100, 7.0), (133, 1204, 20.0), (600, 9000, 2.0)])
def test_forced_overlap_save_row_pass(gb, refcheck, rows, cols, sigma):
    img = go.synth_u8((rows, cols, 4), rows + cols)
    out, inter, launches = _blur_device(gb, img, sigma, {"GBLUR_ROW_OLS": "1", "GBLUR_COL_OLS": "0"})
    assert launches == 3
    _check(out, inter, img, sigma, refcheck, f"{rows}x{cols} sigma={sigma} forced row")


@pytest.mark.parametrize("rows,cols,sigma", [(1500, 4100, 7.0), (160, 7680, 20.0)])
def test_forced_overlap_save_both_passes(gb, refcheck, rows, cols, sigma):
    img = go.synth_u8((rows, cols, 4), rows * 3 + cols)
    out, inter, launches = _blur_device(gb, img, sigma, {"GBLUR_ROW_OLS": "1", "GBLUR_COL_OLS": "1"})
    assert launches == 5
    _check(out, inter, img, sigma, refcheck, f"{rows}x{cols} sigma={sigma} forced row + column")


def test_overlap_save_adversarial_patterns(gb, refcheck):
    """alternating columns / rows drive the reference's Nyquist terms to their maximum"""
    rows, cols, sigma = 900, 4200, 20.0
    for pat in ("altcols", "altrows"):
        img = np.zeros((rows, cols, 4), np.uint8)
        if pat == "altcols":
            img[:, ::2] = 255
        else:
            img[::2] = 255
        out, inter, _ = _blur_device(gb, img, sigma, {"GBLUR_ROW_OLS": "1", "GBLUR_COL_OLS": "1"})
        want = refcheck.blur(img, sigma, True)
        assert np.abs(out.astype(np.int16) - want.astype(np.int16)).max() <= 1, pat


@pytest.mark.parametrize("rows,cols,sigma", [(1500, 4100, 7.0), (160, 7680, 20.0), (2100, 1028, 0.4)])
def test_rgba_alpha_passed_through_on_the_overlap_save_passes(gb, refcheck, rows, cols, sigma):
    """RGBA with apply_to_alpha = false (the reference's command-line default, examples/desktop/main.cpp:80): all four
    channels run through the block kernels and the column kernel puts the input's alpha byte back into its output tile
    (csrc/ols_kernels.cu). Alpha must come back bit-identical (tests/test_gaussianblur.cpp:115-118), RGB within 1 LSB,
    out of place and in place; sigma = 0.4 is the one-tap window (pad 0)."""
    import torch

    img = go.synth_u8((rows, cols, 4), rows * 5 + cols)
    saved = {k: os.environ.get(k) for k in ("GBLUR_ROW_OLS", "GBLUR_COL_OLS")}
    os.environ.update({"GBLUR_ROW_OLS": "1", "GBLUR_COL_OLS": "1"})
    try:
        plan = gb.Plan(rows, cols, 4, sigma)
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    d_in = torch.from_numpy(img).cuda()
    d_out = torch.full_like(d_in, 0x5A)
    plan.run_device(d_in, d_out, False)
    torch.cuda.synchronize()
    kinds = plan.last_kernels()
    out = d_out.cpu().numpy()
    inter = plan.dump_intermediate(False)
    plan.run_device(d_in, d_in, False)  # in place: an item reads the alpha bytes it is about to overwrite
    torch.cuda.synchronize()
    out_inplace = d_in.cpu().numpy()
    # a buffer that is not 16-byte aligned keeps the whole-line kernels (TMA tiles need the alignment) and the result
    flat = torch.empty(img.size + 4, dtype=torch.uint8, device="cuda")
    d_off = flat[4:].view(rows, cols, 4)
    d_off.copy_(torch.from_numpy(img))
    plan.run_device(d_off, d_off, False)
    torch.cuda.synchronize()
    kinds_off = plan.last_kernels()
    out_off = d_off.cpu().numpy()
    plan.close()
    assert kinds == ("overlap_save_1024", "overlap_save_1024"), kinds
    assert kinds_off[1] != "overlap_save_1024", kinds_off
    want = refcheck.blur(img, sigma, False)
    for name, got in (("out of place", out), ("in place", out_inplace), ("unaligned", out_off)):
        assert np.array_equal(got[..., 3], img[..., 3]), f"{name}: alpha changed"
        d = np.abs(got.astype(np.int16) - want.astype(np.int16))
        assert d.max() <= 1, f"{name}: max |delta| = {d.max()} LSB"
    assert np.array_equal(out, out_inplace)
    got2 = inter.reshape(rows, cols, 3)[..., 2]
    assert np.abs(got2 - refcheck.rowpass(img, 2, sigma)).max() < INTER_TOL


def test_column_kernel_alone_stage_reuses_the_terms(gb):
    """gblur_run_stage_device(stages = 4), the hook bench.py times the column block kernel with: after a full run on the
    same image it launches ONE kernel (the column terms of the previous run are reused) and leaves the same bytes"""
    import torch

    img = go.synth_u8((1800, 4100, 4), 0x4321)
    saved = {k: os.environ.get(k) for k in ("GBLUR_ROW_OLS", "GBLUR_COL_OLS")}
    os.environ.update({"GBLUR_ROW_OLS": "1", "GBLUR_COL_OLS": "1"})
    try:
        plan = gb.Plan(1800, 4100, 4, 9.0)
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    d_in = torch.from_numpy(img).cuda()
    d_out = torch.empty_like(d_in)
    plan.run_device(d_in, d_out, True)
    torch.cuda.synchronize()
    want = d_out.cpu().numpy()
    d_out2 = torch.zeros_like(d_in)
    before = gb.launch_count()
    plan.run_stage(4, d_in, d_out2, True)
    torch.cuda.synchronize()
    launches = gb.launch_count() - before
    plan.close()
    assert launches == 1
    assert np.array_equal(d_out2.cpu().numpy(), want)
