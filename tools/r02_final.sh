#!/bin/bash
# Round-end pass on one GPU: whole GPU suite (every test under a time limit), smoke, the default bench line with the CPU leg,
# the reference arm, the small-batch latency, and the ncu launch list of the default command.  -> gpurun_out/<tag>_*
mkdir -p gpurun_out
tag=${1:-r02f}
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/${tag}_pytest_gpu.log 2>&1; grep -n "^E " gpurun_out/${tag}_pytest_gpu.log | cut -c1-300 | head -20; tail -4 gpurun_out/${tag}_pytest_gpu.log
python __graft_entry__.py smoke 2>&1 | tail -2 | tee gpurun_out/${tag}_smoke.log
t0=$(date +%s)
timeout 900 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$? $(( $(date +%s) - t0 )) s"; tail -2 gpurun_out/${tag}_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/${tag}_reference.json 2>gpurun_out/${tag}_reference.err; tail -c 600 gpurun_out/${tag}_reference.json
timeout 300 python tools/small_batch_latency.py 2>&1 | tail -6 | tee gpurun_out/${tag}_small_batch_latency.log
python tools/show_bench.py gpurun_out/${tag}_bench.json
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 0 --sub none --no-cpu-baseline"
$CMD > gpurun_out/${tag}_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/${tag}_ncu_launches.csv $CMD > gpurun_out/${tag}_ncu_list.log 2>&1
echo "launch list rc=$?"
