#!/bin/bash
# Latency of a small batch interval with the single-pass CSR kernel and with the two-pass kernels.
mkdir -p gpurun_out
echo "narrow:"; python tools/small_batch_latency.py 2>&1 | tail -4 | tee gpurun_out/lat_narrow.log
echo "two-pass:"; TWTML_NO_NARROW=1 python tools/small_batch_latency.py 2>&1 | tail -4 | tee gpurun_out/lat_twopass.log
