#!/bin/bash
# Round-end pass on one GPU: whole GPU suite, smoke, the default bench line (with the CPU leg), the reference arm,
# and the other workloads' lines.  Logs and JSON lines -> gpurun_out/final_*.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/final_pytest.log; tail -3 gpurun_out/final_pytest.log
python __graft_entry__.py smoke 2>&1 | tail -2 | tee gpurun_out/final_smoke.log
python bench.py 2>&1 | tail -1 > gpurun_out/final_bench_dense_1m.json
python bench.py --impl reference --steps 2 --warmup 1 2>&1 | tail -1 > gpurun_out/final_bench_reference.json
for w in tweets_1m tweets_1m_wide csr_2m dense_d128; do
  python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/final_bench_$w.json
done
for f in gpurun_out/final_bench_*.json; do echo "$f $(grep -o '"value": [0-9.]*' $f | head -3 | tr '\n' ' ') $(grep -o '"frac": [0-9.]*' $f | head -1)"; done
