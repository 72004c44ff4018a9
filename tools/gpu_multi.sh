#!/bin/bash
# Multi-GPU pass (gpurun --gpus N): 1-vs-2-GPU bit-identity + oracle parity under torchrun, then the
# bench at N = 1, 2, (4, 8).  Logs -> gpurun_out/.
mkdir -p gpurun_out
tag=${1:-mg}
NG=$(nvidia-smi -L | wc -l)
echo "gpus: $NG" | tee gpurun_out/${tag}_env.log
python -m pytest tests/test_multi_rank.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/${tag}_pytest.log
for n in 1 2 4 8; do
  [ $n -le $NG ] || continue
  if [ $n -eq 1 ]; then
    python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | tee gpurun_out/${tag}_bench_n$n.log
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29611 \
      bench.py --gpus $n --steps 5 --warmup 3 2>&1 | tail -1 | tee gpurun_out/${tag}_bench_n$n.log
  fi
done
