#!/usr/bin/env python
"""Short command for `ncu --set full`: three fused runs of one workload (row pass + column pass
per run), nothing else on the GPU. Usage: python tools/prof_cmd.py [workload] [runs]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libgaussianblur_b200 as gb  # noqa: E402
from libgaussianblur_b200.synth import synth_u8  # noqa: E402

WL = {"8k": (4320, 7680, 4, 20.0, True), "1080p": (1080, 1920, 3, 5.0, False),
      "c4row": (4096, 34000, 1, 50.0, False), "c4col": (34000, 4096, 1, 50.0, False)}


def main():
    rows, cols, ch, sigma, alpha = WL[sys.argv[1] if len(sys.argv) > 1 else "8k"]
    runs = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    d_in = torch.from_numpy(synth_u8((rows, cols, ch), 0x1235)).cuda()
    print(gb.host_plan_sizes(rows, cols, sigma))
    d_out = torch.empty_like(d_in)
    plan = gb.Plan(rows, cols, ch, sigma)
    for _ in range(runs):
        plan.run_device(d_in, d_out, alpha)
    torch.cuda.synchronize()
    plan.close()
    print("ok", gb.launch_count())


if __name__ == "__main__":
    main()
