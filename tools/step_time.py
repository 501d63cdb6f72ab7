#!/usr/bin/env python
"""ms per fused 8K RGBA step (device-resident, CUDA events, back to back) — for A/B runs of environment knobs."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import libgaussianblur_b200 as gb
from libgaussianblur_b200.synth import synth_u8

rows, cols, ch, sigma = (int(os.environ.get("STEP_ROWS", "4320")), int(os.environ.get("STEP_COLS", "7680")),
                          int(os.environ.get("STEP_CH", "4")),
                          float(os.environ.get("STEP_SIGMA", "20")))
alpha = os.environ.get("STEP_ALPHA", "1") != "0"
plan = gb.Plan(rows, cols, ch, sigma)
d_in = torch.from_numpy(synth_u8((rows, cols, ch), 0x1235)).cuda()
if ch == 1:
    alpha = False
d_out = torch.empty_like(d_in)
for _ in range(10):
    plan.run_device(d_in, d_out, alpha)
torch.cuda.synchronize()
best = []
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        plan.run_device(d_in, d_out, alpha)
    e1.record()
    torch.cuda.synchronize()
    best.append(e0.elapsed_time(e1) / 200)
print(json.dumps({"shape": [rows, cols, ch], "sigma": sigma, "apply_to_alpha": alpha, "knobs": {k: v for k, v in os.environ.items() if k.startswith("GBLUR_")}, "ms_per_step": best,
                  "kernels": plan.last_kernels()}))
