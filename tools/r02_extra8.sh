#!/bin/bash
# Leftover lines on the 8-GPU box: configs[0] on the GPU at 8 ranks, and the skewed wide-hash tweets on 1.
mkdir -p gpurun_out
tag=${1:-r02n}
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29650 bench.py --gpus 8 --workload tweets_1m --steps 5 --warmup 3 --e2e-steps 4 --sub none 2>gpurun_out/${tag}_tw8.err | tail -1 > gpurun_out/${tag}_scale_tweets_1m_n8.json
timeout 200 python bench.py --workload tweets_1m_wide --steps 5 --warmup 3 --e2e-steps 4 --sub none --no-cpu-baseline 2>gpurun_out/${tag}_tww.err | tail -1 > gpurun_out/${tag}_tweets_1m_wide_n1.json
python tools/show_bench.py gpurun_out/${tag}_scale_tweets_1m_n8.json; python tools/show_bench.py gpurun_out/${tag}_tweets_1m_wide_n1.json
