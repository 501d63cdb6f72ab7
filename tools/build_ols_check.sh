#!/bin/bash
# builds tools/ols_gpu_check (travels to the GPU box with the snapshot; git-ignored)
set -e
cd "$(dirname "$0")/.."
/usr/local/cuda/bin/nvcc -std=c++17 -O2 -cudart static -I include tools/ols_gpu_check.cu \
  -L libgaussianblur_b200/lib -lgblur_b200 -Xlinker -rpath -Xlinker '$ORIGIN/../libgaussianblur_b200/lib' -o tools/ols_gpu_check
