#!/bin/bash
# Short round-end pass on one GPU: whole GPU suite, smoke, the default bench line (CPU leg included).
mkdir -p gpurun_out
tag=${1:-r02q}
t0=$(date +%s)
timeout 600 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/${tag}_pytest_gpu.log 2>&1; grep -n "^E " gpurun_out/${tag}_pytest_gpu.log | cut -c1-300 | head -20; tail -3 gpurun_out/${tag}_pytest_gpu.log
echo "tests $(( $(date +%s) - t0 )) s"
python __graft_entry__.py smoke 2>&1 | tail -2 | tee gpurun_out/${tag}_smoke.log
t0=$(date +%s)
timeout 400 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$? $(( $(date +%s) - t0 )) s"; tail -2 gpurun_out/${tag}_bench.err
python tools/show_bench.py gpurun_out/${tag}_bench.json
