"""gblur_colsum_noalloc (include/gblur_capi.h): the entry point that allocates nothing — the weighted column sums of the
column pass's Nyquist term accumulated in the library's file-scope static __device__ scratch instead of a plan buffer.
No counterpart in the reference, so the check is against the existing
entry point that computes the same integers, gblur_halo_colsum_device on a one-rank halo plan, and against the exact
integer sums in numpy. Integer work: the bar is bit-exact. Two calls in a row on different images show that the shared
scratch is cleared between calls. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _numpy_colsum(img2d, pad):
    rows = img2d.shape[0]
    r = np.arange(rows)
    m = 1 + ((r >= 1) & (r <= pad)).astype(np.int64) + ((r >= rows - 1 - pad) & (r <= rows - 2)).astype(np.int64)
    w = np.where((pad + r) & 1, -m, m)
    return (w[:, None] * img2d.astype(np.int64)).sum(axis=0)


def _existing_entry_point(gb, torch, d_img, rows, cols, ch, sigma):
    plan = gb.HaloPlan(rows, cols, ch, sigma, 1, 0)
    d_u = torch.zeros(cols * ch, dtype=torch.int32, device="cuda")
    plan.colsum(d_img, d_u)
    torch.cuda.synchronize()
    pad = plan.pad
    plan.close()
    return d_u.cpu().numpy(), pad


@pytest.mark.parametrize("rows,cols,ch,sigma", [(3000, 1040, 4, 20.0), (4100, 1024, 1, 20.0)])
def test_colsum_noalloc_matches_existing_entry_point_twice_in_a_row(gb, rows, cols, ch, sigma):
    import torch

    from libgaussianblur_b200.synth import synth_u8

    L = gb.lib()
    imgs = [synth_u8((rows, cols * ch), 0x51 + k) for k in range(2)]
    imgs[1] = 255 - imgs[1]  # a different image with different sums: stale scratch contents would show
    d_imgs = [torch.from_numpy(i).cuda() for i in imgs]
    want = [_existing_entry_point(gb, torch, d, rows, cols, ch, sigma) for d in d_imgs]
    d_out = [torch.full((cols * ch,), -7, dtype=torch.int32, device="cuda") for _ in range(2)]
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()
    free_before = None
    for k in range(2):  # two calls in a row, through the C ABI, same stream (calls that overlap on the GPU would race)
        if k == 1:
            free_before = torch.cuda.mem_get_info()[0]
        rc = L.gblur_colsum_noalloc(0, C.c_void_p(d_imgs[k].data_ptr()), rows, cols, ch, C.c_float(sigma),
                                    C.c_void_p(d_out[k].data_ptr()), C.c_void_p(stream.cuda_stream))
        assert rc == 0, L.gblur_last_error().decode()
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before, "the second call changed the free device memory"
    for k in range(2):
        got = d_out[k].cpu().numpy()
        assert np.array_equal(got, want[k][0]), f"call {k}: differs from gblur_halo_colsum_device"
        assert np.array_equal(got.astype(np.int64), _numpy_colsum(imgs[k], want[k][1])), f"call {k}: differs from numpy"
    assert not np.array_equal(want[0][0], want[1][0])


def test_colsum_noalloc_rejects_rows_beyond_the_static_scratch(gb):
    import torch

    L = gb.lib()
    d = torch.zeros(1024, dtype=torch.uint8, device="cuda")
    o = torch.zeros(1024, dtype=torch.int32, device="cuda")
    cols = gb.NOALLOC_MAX_ROW_BYTES // 4 + 4  # x 4 channels: 16 bytes over the limit
    rc = L.gblur_colsum_noalloc(0, C.c_void_p(d.data_ptr()), 4096, cols, 4, C.c_float(5.0), C.c_void_p(o.data_ptr()), None)
    assert rc == 3  # GBLUR_E_UNSUPPORTED
    msg = L.gblur_last_error().decode()
    assert "GBLUR_NOALLOC_MAX_ROW_BYTES" in msg and str(gb.NOALLOC_MAX_ROW_BYTES) in msg
    # and the widest row that fits is accepted
    rows, cols = 600, gb.NOALLOC_MAX_ROW_BYTES // 4
    img = torch.randint(0, 256, (rows, cols * 4), dtype=torch.uint8, device="cuda")
    out = torch.empty(cols * 4, dtype=torch.int32, device="cuda")
    gb.colsum_noalloc(img, rows, cols, 4, 5.0, out)
    torch.cuda.synchronize()
    pad = gb.host_plan_sizes(rows, cols, 5.0)["pad"]
    assert np.array_equal(out.cpu().numpy().astype(np.int64), _numpy_colsum(img.cpu().numpy(), pad))
