#!/bin/bash
# Wide CSR path: parity tests, then the two wide workloads.
mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests/test_sparse_gpu.py tests/test_fuzz_gpu.py tests/test_narrow_csr_gpu.py -m gpu -x -q 2>&1 | tail -3
for w in csr_2m tweets_1m_wide; do
timeout -k 10 300 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 0 2>&1 | tail -1 > gpurun_out/wide_$w.json
echo "$w $(grep -o '"value": [0-9.]*\|kernel_ms": [0-9.]*' gpurun_out/wide_$w.json | tr '\n' ' ')"
done
