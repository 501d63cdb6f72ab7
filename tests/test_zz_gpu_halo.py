"""Halo sharding (include/gblur_capi.h gblur_halo_*; SURVEY.md 8e third row): row bands for both passes. The ranks are
independent but for `pad` input rows on either side and one integer all-reduce, so the N-rank computation is run here
on ONE GPU, rank after rank (each with its own plan, the column sums added on the device as the all-reduce would), and
the assembled image is compared with the compiled reference. Multi-process plumbing: tests/test_multi_cpu.py (gloo) and
bench.py's band_check (NCCL)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,cols,ch,sigma,ranks", [
    (5000, 4112, 1, 20.0, 3),     # 6 blocks of 894 rows over 3 ranks
    (6000, 2064, 1, 50.0, 4),     # pad 165, 9 blocks over 4 ranks (uneven)
    (3000, 1040, 4, 20.0, 2),     # RGBA, pair-major panel rows
    (2000, 1024, 4, 5.0, 5),      # more ranks than blocks (3 blocks): two ranks idle
    (4100, 1024, 1, 20.0, 1),     # one rank = the whole image through the halo entry points
])
def test_halo_sharded_ranks_vs_reference(gb, refcheck, rows, cols, ch, sigma, ranks):
    import torch

    from libgaussianblur_b200.synth import synth_u8

    img = synth_u8((rows, cols, ch) if ch > 1 else (rows, cols), 0x4A + rows)
    want = refcheck.blur(img, sigma, True)
    d_img = torch.from_numpy(img).cuda().reshape(rows, cols * ch)
    plans = [gb.HaloPlan(rows, cols, ch, sigma, ranks, r) for r in range(ranks)]
    # the partition covers every row exactly once, on block boundaries
    assert plans[0].out_row0 == 0 and plans[-1].out_row1 == rows
    for a, b in zip(plans, plans[1:]):
        assert a.out_row1 == b.out_row0
    d_u = torch.zeros(cols * ch, dtype=torch.int32, device="cuda")
    for p in plans:
        if p.out_row1 > p.out_row0:
            p.colsum(d_img[p.in_row0:p.in_row1], d_u)
    out = torch.empty_like(d_img)
    for p in plans:
        if p.out_row1 > p.out_row0:
            band = d_img[p.in_row0:p.in_row1].contiguous()
            p.run(band, d_u, out[p.out_row0:p.out_row1])
    torch.cuda.synchronize()
    got = out.cpu().numpy().reshape(img.shape)
    for p in plans:
        p.close()
    d = np.abs(got.astype(np.int16) - want.astype(np.int16))
    assert d.max() <= 1, f"max |delta| = {d.max()} LSB at {np.argwhere(d > 1)[:3].tolist()}"


@pytest.mark.parametrize("rows,cols,ch,sigma,ranks", [(6000, 2064, 1, 50.0, 4), (3600, 1040, 4, 20.0, 3)])
def test_halo_rows_above_and_below_in_one_launch(gb, refcheck, rows, cols, ch, sigma, ranks):
    """gblur_halo_rowpass2_device (own rows first, then both halo ranges in ONE launch: what the multi-process binding
    does) against the compiled reference, and against the single launch over all of the rank's rows: identical bytes
    for RGBA (a complex sequence is a channel pair of one row); for one channel a sequence is a PAIR OF ROWS, the split
    launches pair them differently (odd halo heights: a pair even straddles the gap between the two ranges) and fp32
    rounding couples the two lines of a sequence, so there the bar is the usual 1 LSB."""
    import torch

    from libgaussianblur_b200.synth import synth_u8

    img = synth_u8((rows, cols, ch) if ch > 1 else (rows, cols), 0x99 + rows)
    ref = refcheck.blur(img, sigma, True).reshape(rows, cols * ch)
    d_img = torch.from_numpy(img).cuda().reshape(rows, cols * ch)
    plans = [gb.HaloPlan(rows, cols, ch, sigma, ranks, r) for r in range(ranks)]
    d_u = torch.zeros(cols * ch, dtype=torch.int32, device="cuda")
    for p in plans:
        p.colsum(d_img[p.in_row0:p.in_row1], d_u)
    for p in plans:
        band = d_img[p.in_row0:p.in_row1].contiguous()
        n_out = p.out_row1 - p.out_row0
        want = torch.empty((n_out, cols * ch), dtype=torch.uint8, device="cuda")
        p.run(band, d_u, want)
        torch.cuda.synchronize()
        got = torch.empty_like(want)
        p.rowpass(band, p.out_row0, p.out_row1)
        p.rowpass2(band, p.in_row0, p.out_row0, p.out_row1, p.in_row1)
        p.colpass(d_u, got)
        torch.cuda.synchronize()
        g = got.cpu().numpy().astype(np.int16)
        d_ref = np.abs(g - ref[p.out_row0:p.out_row1].astype(np.int16))
        assert d_ref.max() <= 1, f"rank rows [{p.out_row0}, {p.out_row1}): max |delta| vs reference = {d_ref.max()}"
        d_one = np.abs(g - want.cpu().numpy().astype(np.int16))
        if ch == 4:
            assert d_one.max() == 0, f"rank rows [{p.out_row0}, {p.out_row1})"
        else:
            assert d_one.max() <= 1 and (d_one > 0).mean() < 1e-3
    assert any(p.out_row0 > p.in_row0 and p.in_row1 > p.out_row1 for p in plans)
    for p in plans:
        p.close()


def test_halo_partition_host_only(gb):
    for rows, sigma, ranks in ((32768, 50.0, 8), (5000, 20.0, 3), (1000, 5.0, 4)):
        parts = [gb.halo_partition(rows, 4096, sigma, ranks, r) for r in range(ranks)]
        assert parts[0]["out_row0"] == 0 and parts[-1]["out_row1"] == rows
        pad = gb.host_plan_sizes(rows, 4096, sigma)["pad"]
        for p in parts:
            if p["out_row1"] > p["out_row0"]:
                assert p["in_row0"] == max(0, p["out_row0"] - pad)
                assert p["in_row1"] >= min(rows, p["out_row1"] + pad)
