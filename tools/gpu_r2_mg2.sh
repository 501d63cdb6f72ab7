#!/bin/bash
# 2 GPUs: the multi-GPU C ABI from a C++-only program, the C++ API tests and the torchrun tests
mkdir -p gpurun_out/mg2
nvidia-smi -L | wc -l
timeout 1200 python -m pytest tests/test_gpu_mg_capi.py tests/test_gpu_multi.py tests/test_gpu_cpp_api.py -m gpu -x -q -s > gpurun_out/mg2/pytest_mg.log 2>&1; echo "pytest rc=$?"
grep -E "bands|frames|mg check|passed|failed|FAIL" gpurun_out/mg2/pytest_mg.log | head -30
