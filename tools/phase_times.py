#!/usr/bin/env python
"""Row-pass / column-pass split of one workload, timed with CUDA events on the launching stream
(gblur_run_stage_device: the same kernels gblur_run_batch_device launches).
Usage: python tools/phase_times.py rows cols channels sigma alpha frames [reps]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libgaussianblur_b200 as gb  # noqa: E402
from libgaussianblur_b200.synth import synth_u8  # noqa: E402


def main():
    rows, cols, ch = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    sigma, alpha, frames = float(sys.argv[4]), bool(int(sys.argv[5])), int(sys.argv[6])
    reps = int(sys.argv[7]) if len(sys.argv) > 7 else 10
    d_in = torch.from_numpy(synth_u8((frames, rows, cols, ch), 0x1235)).cuda()
    d_out = torch.empty_like(d_in)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    plan = gb.Plan(rows, cols, ch, sigma, max_frames=frames)
    out = {"shape": [frames, rows, cols, ch], "sigma": sigma, "radices": [plan.radices(0), plan.radices(1)]}
    for name, stages in (("row", 1), ("col", 2), ("both", 3)):
        ts = []
        for i in range(reps + 3):
            flush.fill_(i & 0xFF)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            plan.run_stage(stages, d_in, d_out, alpha, n_frames=frames)
            e1.record()
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(e0.elapsed_time(e1))
        ts.sort()
        out[name + "_ms"] = round(ts[len(ts) // 2], 4)
    px = frames * rows * cols
    out["gpx_s_both"] = round(px / out["both_ms"] / 1e6, 2)
    print(json.dumps(out))
    plan.close()


if __name__ == "__main__":
    main()
