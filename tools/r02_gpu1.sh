#!/bin/bash
# Round-2 one-GPU pass: host topology, GPU test suite, smoke, the default bench line (dense_16m + sub-records + CPU leg),
# the reference arm.  Logs and JSON lines -> gpurun_out/<tag>_*.
mkdir -p gpurun_out
tag=${1:-r02a}
{ nvidia-smi --query-gpu=name,memory.total,pci.bus_id --format=csv; nproc; free -g | head -2; lscpu | grep -i -E "numa|socket|model name"; nvidia-smi topo -m; } > gpurun_out/${tag}_env.log 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/${tag}_pytest.log; tail -15 gpurun_out/${tag}_pytest.log
python __graft_entry__.py smoke 2>&1 | tail -5 | tee gpurun_out/${tag}_smoke.log
timeout 900 python bench.py ${BENCH_ARGS} > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 3000 gpurun_out/${tag}_bench.json; tail -5 gpurun_out/${tag}_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/${tag}_reference.json 2>&1; tail -c 800 gpurun_out/${tag}_reference.json
