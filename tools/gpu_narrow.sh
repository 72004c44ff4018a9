#!/bin/bash
# Narrow-CSR kernel: parity tests on the CSR paths, then the tweet workload with the single-pass kernel and with the
# two-pass kernels (TWTML_NO_NARROW=1).  Logs -> gpurun_out/.
mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests/test_sparse_gpu.py tests/test_narrow_csr_gpu.py tests/test_fuzz_gpu.py tests/test_upstream_acceptance.py -m gpu -x -q 2>&1 | tail -25 > gpurun_out/narrow_pytest.log; tail -4 gpurun_out/narrow_pytest.log
for off in "" 1; do
  [ -n "$off" ] && export TWTML_NO_NARROW=1 || unset TWTML_NO_NARROW
  timeout -k 10 200 python bench.py --workload tweets_1m --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/narrow_bench_${off:-0}.json
  echo "no_narrow=${off:-0} $(grep -o '"value": [0-9.]*\|kernel_ms": [0-9.]*\|ingest_ms_per_step": [0-9.]*' gpurun_out/narrow_bench_${off:-0}.json | tr '\n' ' ')"
done
