#!/bin/bash
mkdir -p gpurun_out
for dbg in 32; do
TWTML_NARROW_DBG=$dbg timeout -k 10 200 python bench.py --workload tweets_1m --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 0 2>&1 | tail -1 > gpurun_out/dbg_$dbg.log
echo "dbg=$dbg $(grep -o '"value": [0-9.]*\|kernel_ms": [0-9.]*' gpurun_out/dbg_$dbg.log | tr '\n' ' ')"
done
