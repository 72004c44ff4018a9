#!/bin/bash
# 2-GPU pass for the narrow CSR kernel: multi-rank parity tests, then the tweet workload at N = 1 and 2.
mkdir -p gpurun_out
tag=${1:-mgn}
timeout -k 10 600 python -m pytest tests/test_multi_rank.py -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/${tag}_pytest.log
timeout -k 10 300 python bench.py --workload tweets_1m --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 0 2>&1 | tail -1 > gpurun_out/${tag}_bench_n1.json
NG=$(nvidia-smi -L | wc -l)
for n in 2 4 8; do
  [ $n -le $NG ] || continue
  timeout -k 10 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --workload tweets_1m --gpus $n --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 0 2>&1 | tail -1 > gpurun_out/${tag}_bench_n$n.json
done
for f in gpurun_out/${tag}_bench_n*.json; do echo "$f $(grep -o '"value": [0-9.]*\|"n_gpus": [0-9]*\|"ms_per_step": [0-9.]*' $f | head -3 | tr '\n' ' ')"; done
