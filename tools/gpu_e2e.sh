#!/bin/bash
# Dense ingest: host-to-device rates (tools/pcie_probe.py), the dense parity tests, the default bench line (e2e leg).
mkdir -p gpurun_out
python tools/pcie_probe.py 2>&1 | tail -8
timeout -k 10 900 python -m pytest tests/test_dense_gpu.py tests/test_fuzz_gpu.py tests/test_cpp_facade.py -m gpu -x -q 2>&1 | tail -3
timeout -k 10 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/e2e_bench.json
grep -o '"value": [0-9.]*\|h2d_GBps_effective": [0-9.]*\|ingest_ms_per_step": [0-9.]*\|compute_ms_per_step": [0-9.]*' gpurun_out/e2e_bench.json | tr '\n' ' '
