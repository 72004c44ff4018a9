#!/bin/bash
# Two-GPU functional pass (gpurun --gpus 2): the >= 2-GPU tests (ranks as processes under torchrun, ranks as threads through
# twtml_create_all, the compiled C++ driver on a device list) and the wide-CSR exchange A/B (peer memory vs NCCL).
mkdir -p gpurun_out
tag=${1:-r02x}
timeout 1500 python -m pytest tests/test_multi_rank.py tests/test_group_gpu.py tests/test_cpp_facade.py -m gpu -x -q 2>&1 | tail -25 > gpurun_out/${tag}_pytest_multi.log; tail -8 gpurun_out/${tag}_pytest_multi.log
tr2() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 "${@:2}"; }
tr2 29611 --workload csr_2m --steps 5 --warmup 3 --e2e-steps 4 --sub none 2>gpurun_out/${tag}_csr_p2p.err | tail -1 > gpurun_out/${tag}_csr_2m_n2_p2p.json
TWTML_NO_P2P_WIDE=1 tr2 29612 --workload csr_2m --steps 5 --warmup 3 --e2e-steps 0 --sub none 2>gpurun_out/${tag}_csr_nccl.err | tail -1 > gpurun_out/${tag}_csr_2m_n2_nccl.json
tr2 29613 --workload dense_1m --steps 5 --warmup 3 --e2e-steps 4 --sub none 2>gpurun_out/${tag}_d1m.err | tail -1 > gpurun_out/${tag}_dense_1m_n2.json
tr2 29614 --workload tweets_1m --steps 5 --warmup 3 --e2e-steps 4 --sub none 2>gpurun_out/${tag}_tw.err | tail -1 > gpurun_out/${tag}_tweets_1m_n2.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29700 tools/pcie_probe_multi.py 2>/dev/null | tail -1 > gpurun_out/${tag}_pcie_n2.json; cat gpurun_out/${tag}_pcie_n2.json
for f in gpurun_out/${tag}_*_n2*.json; do echo "$f $(grep -o '"value": [0-9.e+]*' $f | head -2 | tr '\n' ' ') $(grep -o '"exchange": "[^"]*"' $f | tail -1)"; done
tail -3 gpurun_out/${tag}_*.err 2>/dev/null | tail -20
