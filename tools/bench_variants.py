#!/usr/bin/env python
"""Tuning harness (GPU box): time the row / column kernels of the fused run for each compile-time
variant of the headline FFT lengths (static_shapes.h; selected with GBLUR_VARIANT_<N>=<id>) and
check every variant against the default one (<= 1 LSB).

    python tools/bench_variants.py [--row 0,4,5,6] [--col 1,7,8,9] [--iters 10]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libgaussianblur_b200 as gb  # noqa: E402
from libgaussianblur_b200.synth import synth_u8  # noqa: E402


def time_ms(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--row", default="0")
    ap.add_argument("--col", default="1")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--rows", type=int, default=4320)
    ap.add_argument("--cols", type=int, default=7680)
    ap.add_argument("--sigma", type=float, default=20.0)
    a = ap.parse_args()
    rows, cols, ch, sigma = a.rows, a.cols, 4, a.sigma
    img = synth_u8((rows, cols, ch), 0x1235)
    d_in = torch.from_numpy(img).cuda()
    d_out = torch.empty_like(d_in)
    sz = gb.host_plan_sizes(rows, cols, sigma)
    n_row, n_col = sz["n_rowpass"], sz["n_colpass"]
    ref = None
    results = []
    for rv in a.row.split(","):
        for cv in a.col.split(","):
            os.environ[f"GBLUR_VARIANT_{n_row}"] = rv
            os.environ[f"GBLUR_VARIANT_{n_col}"] = cv
            try:
                plan = gb.Plan(rows, cols, ch, sigma)
                plan.run_device(d_in, d_out, True)
                torch.cuda.synchronize()
                out = d_out.cpu().numpy()
                if ref is None:
                    ref = out.copy()
                diff = int(np.abs(out.astype(int) - ref.astype(int)).max())
                t_row = time_ms(lambda: plan.run_stage(1, d_in, d_out, True), a.iters)
                t_col = time_ms(lambda: plan.run_stage(2, d_in, d_out, True), a.iters)
                t_all = time_ms(lambda: plan.run_device(d_in, d_out, True), a.iters)
                rec = dict(row=rv, col=cv, row_radices=plan.radices(0), col_radices=plan.radices(1), row_ms=t_row,
                           col_ms=t_col, total_ms=t_all, mpx_s=rows * cols / t_all / 1e3, max_diff_vs_first=diff)
                plan.close()
            except Exception as e:  # noqa: BLE001
                rec = dict(row=rv, col=cv, error=str(e))
            results.append(rec)
            print(json.dumps(rec), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
