#!/bin/bash
# Narrow-CSR: entry order chosen at ingest against bank conflicts, A/B + tests + ncu counters.  Outputs -> gpurun_out/<tag>_*.
mkdir -p gpurun_out
tag=${1:-r02e}
run() { name=$1; shift; env "$@" timeout 600 python bench.py --workload tweets_1m --steps 10 --warmup 3 --e2e-steps ${E2E:-0} --sub none --no-cpu-baseline > gpurun_out/${tag}_tw_$name.json 2>gpurun_out/${tag}_tw_$name.err; echo "$name $(grep -o '"value": [0-9.e+]*' gpurun_out/${tag}_tw_$name.json | head -2 | tr '\n' ' ') $(grep -o '"kernel_ms": [0-9.e+-]*' gpurun_out/${tag}_tw_$name.json) $(grep -o '"ms_per_step": [0-9.e+-]*' gpurun_out/${tag}_tw_$name.json | head -1)"; tail -2 gpurun_out/${tag}_tw_$name.err; }
E2E=8 run permuted A=1
run canonical TWTML_NO_NARROW_PERMUTE=1
run permuted_rows2s8 TWTML_NARROW_ROWS2=1 TWTML_NARROW_STAGES=8
timeout 900 python -m pytest tests/test_narrow_csr_gpu.py tests/test_sparse_gpu.py tests/test_parity_edges_gpu.py tests/test_ingest_async_gpu.py tests/test_fuzz_gpu.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/${tag}_pytest.log
timeout 300 python tools/stress_narrow.py 40 13 2>&1 | tail -2 | tee gpurun_out/${tag}_stress.log
TCMD="python bench.py --workload tweets_1m --steps 1 --warmup 3 --e2e-steps 0 --sub none --no-cpu-baseline"
$TCMD > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_csr_narrow -s 20 -c 1 -o gpurun_out/${tag}_csr_narrow_permuted -f $TCMD > gpurun_out/${tag}_ncu_narrow.log 2>&1
echo "ncu rc=$?"
