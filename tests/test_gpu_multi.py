"""Two-GPU test of the band-sharded path (NCCL all-to-all); skipped on boxes with one GPU."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.environ["GBLUR_ROOT"])
from libgaussianblur_b200.multi import BandShardedBlur, FrameShardedBlur, shard_range
from oracle import gblur_oracle as go
from oracle import ref_lib
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
for ch, alpha, fused in ((1, False, True), (3, False, True), (4, False, True), (4, True, True), (1, False, False), (3, False, False), (4, False, False)):
    rows, cols, sigma = 250, 384, 6.0
    img = go.synth_u8((rows, cols, ch) if ch > 1 else (rows, cols), 5 + ch)
    want = (ref_lib.gaussianblur(img, sigma, alpha) if ch > 1 else ref_lib.blur_plane_u8(img, sigma)) if ref_lib.available() else go.gaussianblur(img, sigma, alpha)
    b = BandShardedBlur.cuda(rows, cols, ch, sigma, alpha, device=local, fused=fused)
    assert (b.fused_run is not None) == fused, b.exchange_kind
    r0, r1 = b.my_rows
    out = b.run(torch.from_numpy(img.reshape(rows, cols, ch)[r0:r1].copy()).cuda(), gather_rows=True)
    torch.cuda.synchronize()
    d = np.abs(out.cpu().numpy().astype(int) - want.reshape(rows, cols, ch)[r0:r1].astype(int)).max()
    assert d <= 1, (ch, alpha, fused, d)
    b.close()
for ch, alpha in ((1, False), (4, False)):
    rows, cols, sigma = 250, 384, 6.0
    img = go.synth_u8((rows, cols, ch) if ch > 1 else (rows, cols), 9 + ch)
    want = (ref_lib.gaussianblur(img, sigma, alpha) if ch > 1 else ref_lib.blur_plane_u8(img, sigma)) if ref_lib.available() else go.gaussianblur(img, sigma, alpha)
    b = BandShardedBlur.cuda(rows, cols, ch, sigma, alpha, device=local, fused=False, pipeline_chunks=3)
    assert "pipelined" in b.exchange_kind
    r0, r1 = b.my_rows
    out = b.run(torch.from_numpy(img.reshape(rows, cols, ch)[r0:r1].copy()).cuda(), gather_rows=True)
    torch.cuda.synchronize()
    d = np.abs(out.cpu().numpy().astype(int) - want.reshape(rows, cols, ch)[r0:r1].astype(int)).max()
    assert d <= 1, ("pipelined", ch, d)
    b.close()
for ch, alpha, chunks, sigma, kind in ((1, False, 3, 6.0, "staged"), (3, False, 1, 6.0, "staged"), (4, True, 4, 6.0, "staged"),
                                       (1, False, 2, 40.0, "staged"), (1, False, 4, 40.0, "hybrid"), (3, False, 5, 6.0, "hybrid")):
    rows, cols = 250, 384  # sigma=40: pad 132 > the 125-row bands, so both ranks write reflect_101 mirror slots
    img = go.synth_u8((rows, cols, ch) if ch > 1 else (rows, cols), 19 + ch)
    want = (ref_lib.gaussianblur(img, sigma, alpha) if ch > 1 else ref_lib.blur_plane_u8(img, sigma)) if ref_lib.available() else go.gaussianblur(img, sigma, alpha)
    b = BandShardedBlur.cuda(rows, cols, ch, sigma, alpha, device=local, fused=kind, pipeline_chunks=chunks)
    assert kind in b.exchange_kind
    r0, r1 = b.my_rows
    band = torch.from_numpy(img.reshape(rows, cols, ch)[r0:r1].copy()).cuda()
    for rep in range(2):  # the second image reuses the staging and panel buffers
        out = b.run(band, gather_rows=True)
        torch.cuda.synchronize()
        d = np.abs(out.cpu().numpy().astype(int) - want.reshape(rows, cols, ch)[r0:r1].astype(int)).max()
        assert d <= 1, (kind, ch, chunks, rep, d)
    b.close()
fs = FrameShardedBlur(6, 135, 240, 3, 5.0, device=local, chunk_frames=2)
f0, f1 = fs.frames
frames = np.stack([go.synth_u8((135, 240, 3), 100 + f) for f in range(6)])
o = fs.run(torch.from_numpy(frames[f0:f1]).cuda(), None, False)
torch.cuda.synchronize()
for k, f in enumerate(range(f0, f1)):
    w = ref_lib.gaussianblur(frames[f], 5.0, False) if ref_lib.available() else go.gaussianblur(frames[f], 5.0, False)
    assert np.abs(o[k].cpu().numpy().astype(int) - w.astype(int)).max() <= 1
fs.close()
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
'''


def test_two_gpu_band_and_frame_sharding(tmp_path):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, GBLUR_ROOT=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29541", str(script)],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
