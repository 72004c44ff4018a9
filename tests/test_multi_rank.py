"""Multi-rank host logic.  CPU part: world_size-2 gloo process group exercising the rendezvous that
feeds twtml_create_ranked (id exchange, row ranges, loud failure without a GPU) and bench.py's
reference arm under torchrun.  GPU part (needs >= 2 GPUs, skipped otherwise): tools/mgpu_check.py
under torchrun, and the 1-GPU vs 2-GPU bit-identity of the trained weights."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, %r)
import torch.distributed as dist
import twtml_b200 as tw
from twtml_b200.dist import exchange_nccl_id, row_range, ranked_context
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
uid = exchange_nccl_id(lambda: bytes(range(128)))
assert uid == bytes(range(128))
spans = [row_range(1001, r, world) for r in range(world)]
assert spans[0][0] == 0 and spans[-1][1] == 1001 and all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
try:
    real = exchange_nccl_id()          # the library's own generator (libnccl via dlopen) works without a GPU
    ok_id = len(real) == 128
except tw.TwtmlError as e:
    ok_id = e.code == tw._lib.E_NCCL   # no libnccl on this host: reported, not hidden
assert ok_id
try:
    tw.Context(0, rank, world, uid)
    print("UNEXPECTED context without a GPU")
except tw.TwtmlError as e:
    assert e.code == tw._lib.E_CUDA and "no CPU fallback" in e.message
    print("RANK %%d OK" %% rank)
dist.barrier()
dist.destroy_process_group()
''' % ROOT


def _torchrun(nproc, script_args, env=None, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 2000)] + script_args
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=dict(os.environ, **(env or {})), cwd=ROOT)


def test_gloo_world2_rendezvous(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("CPU-side test")
    w = tmp_path / "worker.py"
    w.write_text(WORKER)
    res = _torchrun(2, [str(w)])
    assert res.returncode == 0, res.stdout + res.stderr
    assert "RANK 0 OK" in res.stdout and "RANK 1 OK" in res.stdout


def test_bench_reference_arm_under_torchrun():
    res = _torchrun(2, ["bench.py", "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1"])
    assert res.returncode == 0, res.stdout + res.stderr
    lines = [json.loads(l) for l in res.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1                                      # rank 0 alone prints; the other rank exits 0
    assert lines[0]["impl"] == "reference" and lines[0]["cpu_baseline"]["kind"] == "port" and lines[0]["value"] > 0


@pytest.mark.gpu
def test_two_gpus_match_one_gpu_bit_for_bit(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    outs = {}
    for g in (1, 2):
        out = str(tmp_path / ("w%d.npz" % g))
        res = _torchrun(g, ["tools/mgpu_check.py"], env={"MGPU_OUT": out}, timeout=600)
        assert res.returncode == 0 and "MGPU OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
        outs[g] = np.load(out)
    for k in outs[1].files:
        if k.startswith("csr"):      # CSR batches use one shard per rank: the sums differ in their last bits
            assert np.allclose(outs[1][k], outs[2][k], rtol=1e-9, atol=1e-300), k
            continue
        if not k[0].isdigit():       # "rep..." (iterations run by repeated calls) and "async" (n % 8 == 0): the same bits
            assert np.array_equal(outs[1][k], outs[2][k]), k
            continue
        n = int(k.split("_")[0])
        if n % 8 == 0:
            assert np.array_equal(outs[1][k], outs[2][k]), k       # same shard vectors, same order, same bits
        else:
            assert np.allclose(outs[1][k], outs[2][k], rtol=1e-9, atol=0)


@pytest.mark.gpu
def test_one_process_one_thread_per_gpu():
    """The single-JVM shape: ranks are host threads of one process (tools/mgpu_threads_check.py)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    res = subprocess.run([sys.executable, "tools/mgpu_threads_check.py", "2"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0 and "THREADS OK 2" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
