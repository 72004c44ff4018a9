#!/bin/bash
# Full GPU test pass + smoke + the sparse bench lines.  Logs -> gpurun_out/.
mkdir -p gpurun_out
tag=${1:-chk}
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${tag}_pytest.log; tail -4 gpurun_out/${tag}_pytest.log
python __graft_entry__.py smoke 2>&1 | tail -3 | tee gpurun_out/${tag}_smoke.log
for w in tweets_1m tweets_1m_wide csr_2m; do
python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 4 2>&1 | tail -1 > gpurun_out/${tag}_bench_$w.json
echo "$w $(grep -o '"value": [0-9.]*\|kernel_ms": [0-9.]*\|ms_per_step": [0-9.]*' gpurun_out/${tag}_bench_$w.json | tr '\n' ' ')"
done
