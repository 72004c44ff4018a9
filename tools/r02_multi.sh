#!/bin/bash
# Round-2 pass on the multi-GPU box (gpurun --gpus 8), most valuable lines first: config 4's protocol -- weak scaling at
# 2^24 rows/GPU (d = 1004) and strong scaling of 2^27 x 128 on 8/1/4/2 GPUs --, config 5 at 8 GPUs (peer-memory vs NCCL
# sliced exchange), the host-link probe with all GPUs copying at once, then the >= 2-GPU tests (ranks as processes, ranks as
# threads of one process through twtml_create_all).  Every command has its own time limit.
# JSON lines -> gpurun_out/<tag>_*.json.   usage: r02_multi.sh <tag> [max gpus]
mkdir -p gpurun_out
tag=${1:-r02m}
G=${2:-8}
trun() { n=$1; shift; if [ "$n" = 1 ]; then python "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) "$@"; fi; }
export -f trun
{ nvidia-smi --query-gpu=name,memory.total,pci.bus_id --format=csv; nproc; free -g | head -2; lscpu | grep -i -E "numa|socket|model name"; nvidia-smi topo -m; } > gpurun_out/${tag}_env.log 2>&1
weak() { n=$1; e2e=$2; timeout 300 bash -c "trun $n bench.py --gpus $n --steps 2 --warmup 3 --e2e-steps $e2e --sub none --no-cpu-baseline" 2>gpurun_out/${tag}_weak_n$n.err | tail -1 > gpurun_out/${tag}_scale_weak_dense16m_n$n.json; }
strong() { n=$1; timeout 200 bash -c "trun $n bench.py --gpus $n --workload dense_d128 --rows $((134217728 / n)) --scaling strong --steps 2 --warmup 3 --e2e-steps 0 --sub none --no-cpu-baseline" 2>gpurun_out/${tag}_strong_n$n.err | tail -1 > gpurun_out/${tag}_scale_strong_d128_n$n.json; }
weak $G 2
strong $G
timeout 200 bash -c "trun $G bench.py --gpus $G --workload csr_2m --steps 5 --warmup 3 --e2e-steps 4 --sub none" 2>gpurun_out/${tag}_csr.err | tail -1 > gpurun_out/${tag}_scale_csr_2m_n${G}_p2p.json
TWTML_NO_P2P_WIDE=1 timeout 200 bash -c "trun $G bench.py --gpus $G --workload csr_2m --steps 5 --warmup 3 --e2e-steps 0 --sub none" 2>gpurun_out/${tag}_csr_nccl.err | tail -1 > gpurun_out/${tag}_scale_csr_2m_n${G}_nccl.json
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29700 tools/pcie_probe_multi.py 2>/dev/null | tail -1 > gpurun_out/${tag}_pcie_n$G.json; cat gpurun_out/${tag}_pcie_n$G.json
strong 1
weak 1 0
for n in 4 2; do [ $n -lt $G ] && { weak $n 0; strong $n; }; done
timeout 200 bash -c "trun $G bench.py --gpus $G --workload dense_1m --steps 5 --warmup 3 --e2e-steps 4 --sub none" 2>gpurun_out/${tag}_d1m.err | tail -1 > gpurun_out/${tag}_scale_dense_1m_n$G.json
timeout 240 python -m pytest tests/test_cpp_facade.py tests/test_group_gpu.py tests/test_multi_rank.py -m gpu -q --timeout 100 > gpurun_out/${tag}_pytest_multi.log 2>&1; grep -n "^E " gpurun_out/${tag}_pytest_multi.log | cut -c1-300 | head -20; tail -4 gpurun_out/${tag}_pytest_multi.log
for f in gpurun_out/${tag}_scale_*.json; do echo "$f $(grep -o '"value": [0-9.e+]*' $f | head -2 | tr '\n' ' ') $(grep -o '"weights_sha256_16": "[0-9a-f]*"' $f | head -1)"; done
tail -n 2 gpurun_out/${tag}_*.err 2>/dev/null | grep -v "^$" | tail -20
