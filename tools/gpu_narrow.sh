#!/bin/bash
# Narrow-CSR experiment: parity tests on the CSR paths, then the tweet workload with and without the single-pass kernel.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sparse_gpu.py tests/test_fuzz_gpu.py tests/test_upstream_acceptance.py -m gpu -x -q 2>&1 | tail -25 > gpurun_out/narrow_pytest.log; tail -8 gpurun_out/narrow_pytest.log
timeout 300 python bench.py --workload tweets_1m --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 2 2>&1 | tail -3 | tee gpurun_out/narrow_bench.log
TWTML_NO_NARROW=1 timeout 300 python bench.py --workload tweets_1m --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 2 2>&1 | tail -3 | tee gpurun_out/narrow_bench_off.log
