#!/bin/bash
# Two-GPU test pass only (gpurun --gpus 2); every test under a 150 s limit so that a hang costs minutes, not the call.
mkdir -p gpurun_out
tag=${1:-r02y}
timeout 900 python -m pytest tests/test_group_gpu.py tests/test_cpp_facade.py tests/test_multi_rank.py tests/test_kmeans_gpu.py tests/test_ingest_async_gpu.py -m gpu -q --timeout 150 > gpurun_out/${tag}_pytest_multi.log 2>&1; grep -n "^E " gpurun_out/${tag}_pytest_multi.log | cut -c1-300 | head -40; tail -8 gpurun_out/${tag}_pytest_multi.log
