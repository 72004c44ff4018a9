#!/bin/bash
# Narrow-CSR kernel experiments on tweets_1m (one GPU): ring depth, row pairs, entries kept in registers; then a full ncu
# capture (with source counters) of the default kernel.  Outputs -> gpurun_out/<tag>_*.
mkdir -p gpurun_out
tag=${1:-r02c}
run() { name=$1; shift; env "$@" timeout 600 python bench.py --workload tweets_1m --steps 10 --warmup 3 --e2e-steps ${E2E:-0} --sub none --no-cpu-baseline > gpurun_out/${tag}_tw_$name.json 2>gpurun_out/${tag}_tw_$name.err; echo "$name $(grep -o '"value": [0-9.e+]*' gpurun_out/${tag}_tw_$name.json | head -2 | tr '\n' ' ') $(grep -o '"kernel_ms": [0-9.e+-]*' gpurun_out/${tag}_tw_$name.json)"; }
E2E=8 run default A=1
run stages8 TWTML_NARROW_STAGES=8
run pairs8 TWTML_NARROW_STAGES=8 TWTML_NARROW_PAIRS=1
if [ -f twitter-stream-ml_b200/libtwtml_k4.so ]; then
  run k4 TWTML_LIB=$PWD/twitter-stream-ml_b200/libtwtml_k4.so
  run k4s8 TWTML_LIB=$PWD/twitter-stream-ml_b200/libtwtml_k4.so TWTML_NARROW_STAGES=8
fi
TCMD="python bench.py --workload tweets_1m --steps 1 --warmup 3 --e2e-steps 0 --sub none --no-cpu-baseline"
$TCMD > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_csr_narrow -s 20 -c 1 -o gpurun_out/${tag}_csr_narrow_default -f $TCMD > gpurun_out/${tag}_ncu_narrow.log 2>&1
echo "narrow full rc=$?"
