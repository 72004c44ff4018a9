#!/bin/bash
# ncu evidence for the dense iteration kernel (one GPU).  Each ncu run follows a plain run of the
# same command that exited 0.  Outputs -> gpurun_out/.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 0 ${BENCH_ARGS}"
tag=${1:-prof}
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_list.log 2>&1
$CMD > gpurun_out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:${KREGEX:-k_dense_iter} -s ${KSKIP:-30} -c ${KCOUNT:-3} -o gpurun_out/${tag}_dense_iter -f $CMD > gpurun_out/${tag}_ncu_full.log 2>&1
tail -2 gpurun_out/${tag}_plain.log; tail -3 gpurun_out/${tag}_ncu_full.log; ls -la gpurun_out/
