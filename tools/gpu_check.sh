#!/bin/bash
# One GPU-box pass: environment probe, GPU test suite, smoke, short bench.  Logs -> gpurun_out/.
mkdir -p gpurun_out
tag=${1:-run}
nvidia-smi --query-gpu=name,memory.total --format=csv | tee gpurun_out/${tag}_env.log
{ nproc; free -g | head -2; java -version 2>&1 | head -1; } >> gpurun_out/${tag}_env.log 2>&1
python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/${tag}_pytest.log; tail -15 gpurun_out/${tag}_pytest.log
python __graft_entry__.py smoke 2>&1 | tail -5 | tee gpurun_out/${tag}_smoke.log
python bench.py --steps ${2:-5} --warmup 3 2>&1 | tail -5 | tee gpurun_out/${tag}_bench.log
