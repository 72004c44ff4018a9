"""The C++ host facade (include/twtml.hpp): compiles as C++17 against the C ABI (CPU check), and on a
GPU replays the reference's driver loop from a compiled host program, checked against the oracle."""
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path):
    from twtml_b200 import _lib
    so = _lib.build()
    exe = str(tmp_path / "driver_loop")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "driver_loop.cpp"), "-o", exe, so,
                           "-Wl,-rpath," + os.path.dirname(so)])
    return exe


def test_cpp_facade_compiles_and_fails_loudly_without_gpu(tmp_path):
    import torch
    exe = _build(tmp_path)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    res = subprocess.run([exe], capture_output=True, text=True)
    assert res.returncode == 1 and "no CPU fallback" in res.stdout


@pytest.mark.gpu
def test_cpp_driver_loop_matches_oracle(tmp_path, oracle):
    exe = _build(tmp_path)
    d = 1004
    files, data = [], []
    for t, n in enumerate([900, 0, 1300]):
        X, y = oracle.synth_dense(n, d, seed=300 + t) if n else (np.zeros((0, d)), np.zeros(0))
        y = np.round(y * 300 + 400)
        p = tmp_path / ("b%d.bin" % t)
        with open(p, "wb") as f:
            f.write(np.array([n, d], np.int64).tobytes()); f.write(np.ascontiguousarray(X).tobytes()); f.write(y.tobytes())
        files.append(str(p)); data.append((X, y))
    res = subprocess.run([exe] + files, capture_output=True, text=True)
    assert res.returncode == 0, res.stdout + res.stderr
    out = json.loads(res.stdout)
    w_ref = np.zeros(d)
    for (X, y), b in zip(data, out["batches"]):
        assert b["count"] == len(y)
        if len(y) == 0:
            assert "mse" not in b                      # rdd.isEmpty: both output ops skip the interval
            continue
        s_ref, _, rnd = oracle.predict_stats(y, d, w_ref, X=X)
        assert (b["mse"], b["real_stdev"], b["pred_stdev"], b["pred0"]) == (s_ref["mse"], s_ref["real_stdev"], s_ref["pred_stdev"], rnd[0])
        w_ref, h_ref = oracle.sgd(y, d, w_ref, X=X, step_size=0.005, num_iterations=50)
        assert b["iters"] == len(h_ref)
    w = np.array(out["weights"])
    assert np.max(np.abs(w - w_ref)) / np.max(np.abs(w_ref)) < 1e-6


@pytest.mark.gpu
def test_cpp_driver_loop_on_all_gpus_through_one_handle(tmp_path, oracle):
    """The same compiled driver, one process, every visible GPU behind one twtml_create_all handle: with row counts that
    are multiples of 8 the weights after three intervals are the single-GPU run's, bit for bit, and the per-interval
    numbers are equal."""
    import torch
    g = torch.cuda.device_count()
    if g < 2:
        pytest.skip("needs >= 2 GPUs")
    g = 8 if g >= 8 else (4 if g >= 4 else 2)
    exe = _build(tmp_path)
    d = 1004
    files = []
    for t, n in enumerate([904, 0, 1296]):
        X, y = oracle.synth_dense(n, d, seed=700 + t) if n else (np.zeros((0, d)), np.zeros(0))
        y = np.round(y * 300 + 400)
        p = tmp_path / ("g%d.bin" % t)
        with open(p, "wb") as f:
            f.write(np.array([n, d], np.int64).tobytes()); f.write(np.ascontiguousarray(X).tobytes()); f.write(y.tobytes())
        files.append(str(p))
    one = subprocess.run([exe] + files, capture_output=True, text=True)
    many = subprocess.run([exe, "--devices", ",".join(str(i) for i in range(g))] + files, capture_output=True, text=True)
    assert one.returncode == 0 and many.returncode == 0, one.stdout + one.stderr + many.stdout + many.stderr
    last_json = lambda out: json.loads([l for l in out.splitlines() if l.startswith("{")][-1])   # libnccl prints its version first
    a, b = last_json(one.stdout), last_json(many.stdout)
    assert a["batches"] == b["batches"]
    assert a["weights"] == b["weights"]                    # printed with %.17g: equal text means equal bits
