#!/bin/bash
# One test file (or pytest selection in $T) on the GPU box.
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest ${T:-tests/test_narrow_csr_gpu.py} -m gpu -x -q 2>&1 | tail -30 | tee gpurun_out/t_pytest.log
