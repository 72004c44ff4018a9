#!/bin/bash
# Eight GPUs: configs[0] and configs[1] with the line-based peer exchange (gpurun --gpus 8).
mkdir -p gpurun_out
tag=${1:-r02p}
trun() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 "${@:2}" 2> gpurun_out/${tag}_$1.err; }
trun 29621 --workload tweets_1m --steps 5 --warmup 3 --sub none --no-cpu-baseline > gpurun_out/${tag}_tweets_1m_n8.json
trun 29622 --workload dense_1m --steps 5 --warmup 3 --sub none --no-cpu-baseline > gpurun_out/${tag}_dense_1m_n8.json
for f in tweets_1m_n8 dense_1m_n8; do echo $f; python tools/show_bench.py gpurun_out/${tag}_$f.json; done
