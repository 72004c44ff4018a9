#!/bin/bash
# Two GPUs, the device-group and multi-rank tests only (short; every test under a time limit).
mkdir -p gpurun_out
timeout 70 python -m pytest tests/test_group_gpu.py tests/test_multi_rank.py -m gpu -q -x --timeout 40 > gpurun_out/r02s_pytest_multi.log 2>&1; grep -n "^E " gpurun_out/r02s_pytest_multi.log | cut -c1-300 | head -20; tail -3 gpurun_out/r02s_pytest_multi.log
