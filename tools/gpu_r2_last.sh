#!/bin/bash
mkdir -p gpurun_out/q
cat > /tmp/c1_mid.py <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, ".")
import libgaussianblur_b200 as gb
from libgaussianblur_b200.synth import synth_u8
from oracle import ref_lib
for rows, cols, sigma in ((4096, 4096, 50.0), (2048, 2064, 10.0), (3000, 4112, 30.0)):
    img = synth_u8((rows, cols), 0x31 + rows)
    want = ref_lib.blur_plane_u8(img, sigma)
    plan = gb.Plan(rows, cols, 1, sigma)
    d = torch.from_numpy(img).cuda().reshape(rows, cols, 1)
    o = torch.empty_like(d)
    plan.run_device(d, o, False); torch.cuda.synchronize()
    k = plan.last_kernels(); plan.close()
    diff = np.abs(o.cpu().numpy().reshape(rows, cols).astype(int) - want.astype(int))
    print(rows, cols, sigma, k, "max", diff.max(), "ndiff", int((diff > 0).sum()))
    assert diff.max() <= 1
print("c1 mid ok")
PY
timeout 300 python /tmp/c1_mid.py 2>&1 | tail -5
for shape in "4096 4096 50" "2048 2048 10"; do
  set -- $shape
  STEP_CH=1 STEP_ROWS=$1 STEP_COLS=$2 STEP_SIGMA=$3 timeout 200 python tools/step_time.py 2>/dev/null | tee -a gpurun_out/q/gates_c1_after.jsonl
done
