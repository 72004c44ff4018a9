#!/bin/bash
# A/B of the shard-vector exchange on N GPUs: peer-memory stores (default) vs NCCL all-gather (TWTML_NO_P2P=1).
mkdir -p gpurun_out
tag=${1:-p2p}
NG=$(nvidia-smi -L | wc -l)
[ -n "$SKIP_TESTS" ] || python -m pytest tests/test_multi_rank.py -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/${tag}_pytest.log
for mode in p2p nccl; do
  if [ $mode = nccl ]; then export TWTML_NO_P2P=1; else unset TWTML_NO_P2P; fi
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29621 \
    bench.py --gpus $NG --steps 10 --warmup 3 --e2e-steps 0 2>&1 | tail -1 | tee gpurun_out/${tag}_bench_${mode}_n$NG.log
done
