#!/bin/bash
# Two-GPU test pass only (gpurun --gpus 2).
mkdir -p gpurun_out
tag=${1:-r02y}
timeout 1500 python -m pytest tests/test_group_gpu.py tests/test_cpp_facade.py tests/test_multi_rank.py -m gpu -q 2>&1 > gpurun_out/${tag}_pytest_multi.log; grep -n "^E " gpurun_out/${tag}_pytest_multi.log | head -60; tail -8 gpurun_out/${tag}_pytest_multi.log
