#!/bin/bash
# Narrow-CSR kernel: parity tests on the CSR paths, then the tweet workload with prefetch distances and without the kernel.
mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests/test_sparse_gpu.py tests/test_fuzz_gpu.py tests/test_upstream_acceptance.py -m gpu -x -q 2>&1 | tail -25 > gpurun_out/narrow_pytest.log; tail -8 gpurun_out/narrow_pytest.log
for ahead in ${AHEADS:-1024 0}; do
TWTML_NARROW_AHEAD=$ahead timeout -k 10 200 python bench.py --workload tweets_1m --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 4 2>&1 | tail -1 > gpurun_out/narrow_bench_$ahead.log
echo "ahead=$ahead $(grep -o '"value": [0-9.]*\|kernel_ms": [0-9.]*\|ingest_ms_per_step": [0-9.]*' gpurun_out/narrow_bench_$ahead.log | tr '\n' ' ')"
done
