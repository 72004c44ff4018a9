#!/bin/bash
# Two GPUs: the multi-rank tests, then configs[0] with the narrow-CSR shard vectors over peer memory (one launch / two
# launches per tail) vs NCCL, and configs[1] as the dense check.
mkdir -p gpurun_out
tag=${1:-r02z}
timeout 600 python -m pytest tests/test_group_gpu.py tests/test_multi_rank.py tests/test_cpp_facade.py -m gpu -q --timeout 150 > gpurun_out/${tag}_pytest_multi.log 2>&1; grep -n "^E " gpurun_out/${tag}_pytest_multi.log | cut -c1-300 | head -40; tail -5 gpurun_out/${tag}_pytest_multi.log
trun() { timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 "${@:2}" 2> gpurun_out/${tag}_$1.err; }
trun 29611 --workload tweets_1m --steps 5 --warmup 3 --sub none --no-cpu-baseline > gpurun_out/${tag}_tweets_n2_p2p.json
TWTML_NO_FUSED_TAIL=1 trun 29612 --workload tweets_1m --steps 5 --warmup 3 --sub none --no-cpu-baseline > gpurun_out/${tag}_tweets_n2_p2p_two_launches.json
TWTML_NO_P2P_NARROW=1 trun 29613 --workload tweets_1m --steps 5 --warmup 3 --sub none --no-cpu-baseline > gpurun_out/${tag}_tweets_n2_nccl.json
trun 29614 --workload dense_1m --steps 5 --warmup 3 --sub none --no-cpu-baseline > gpurun_out/${tag}_dense_1m_n2.json
for f in tweets_n2_p2p tweets_n2_p2p_two_launches tweets_n2_nccl dense_1m_n2; do echo $f; python tools/show_bench.py gpurun_out/${tag}_$f.json; done
