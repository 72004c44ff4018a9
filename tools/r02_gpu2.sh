#!/bin/bash
# Round-2 one-GPU measurement pass: the default bench line (dense_16m + sub-records + CPU leg), a tweets_1m line, and the ncu evidence (launch list of the default command, full-set captures of the two top kernels).
# Every ncu run follows a plain run of the same command that exited 0.  Outputs -> gpurun_out/<tag>_*.
mkdir -p gpurun_out
tag=${1:-r02b}
t0=$(date +%s)
timeout 900 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$? $(( $(date +%s) - t0 )) s"; tail -c 1500 gpurun_out/${tag}_bench.json; tail -3 gpurun_out/${tag}_bench.err
timeout 600 python bench.py --workload tweets_1m --steps 10 --warmup 3 --sub none --no-cpu-baseline > gpurun_out/${tag}_tweets_1m.json 2>gpurun_out/${tag}_tweets.err
# ncu: launch list of the default command (shortened: one timed step), then full-set captures
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 0 --sub none --no-cpu-baseline"
$CMD > gpurun_out/${tag}_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/${tag}_ncu_launches.csv $CMD > gpurun_out/${tag}_ncu_list.log 2>&1
echo "launch list rc=$?"
$CMD > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_dense_iter -s 20 -c 1 -o gpurun_out/${tag}_dense_iter_16m -f $CMD > gpurun_out/${tag}_ncu_dense.log 2>&1
echo "dense full rc=$?"
TCMD="python bench.py --workload tweets_1m --steps 1 --warmup 3 --e2e-steps 0 --sub none --no-cpu-baseline"
$TCMD > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_csr_narrow -s 20 -c 1 -o gpurun_out/${tag}_csr_narrow -f $TCMD > gpurun_out/${tag}_ncu_narrow.log 2>&1
echo "narrow full rc=$?"
ls -la gpurun_out/ | grep ${tag}
