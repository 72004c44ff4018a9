#!/usr/bin/env python3
"""One process, one host thread per GPU (how a single-JVM Spark local[N] executor would drive several GPUs):
each thread owns a ranked context; the micro-batch is row-sharded over them; results must match the CPU oracle
and be identical on every rank.  ctypes releases the GIL inside libtwtml calls, so the ranks really run
concurrently and meet in the library's NCCL all-gather.  Prints 'THREADS OK n'."""
import os
import sys
import threading

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import twtml_b200 as tw
from twtml_b200.dist import row_range
from oracle import oracle_c


def main():
    import torch
    world = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
    n, d, iters = 8 * 1250, 1004, 10
    X, y = oracle_c.synth_dense(n, d, seed=11)
    y = np.round(y * 300 + 400)
    w0 = np.linspace(-0.1, 0.1, d)
    uid = tw.Context.nccl_unique_id()
    out, err = [None] * world, []

    def rank_main(r):
        try:
            ctx = tw.Context(r, r, world, uid)
            lo, hi = row_range(n, r, world)
            m = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(iters).setStepSize(0.1)
                 .setConvergenceTol(0.0).setInitialWeights(tw.Vectors.dense(w0)))
            st, _, _ = m.step(ctx.batch_from_dense(X[lo:hi], y[lo:hi]))
            out[r] = (m.latestModel().weights.toArray(), m.lastLossHistory.copy(), st)
            ctx.close()
        except Exception as e:      # noqa: BLE001
            err.append((r, repr(e)))

    th = [threading.Thread(target=rank_main, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join(300) for t in th]
    assert not err and all(o is not None for o in out), err
    w_ref, h_ref = oracle_c.sgd(y, d, w0, X=X, step_size=0.1, num_iterations=iters, convergence_tol=0.0)
    s_ref, _, _ = oracle_c.predict_stats(y, d, w0, X=X)
    for w, h, st in out:
        assert np.array_equal(w, out[0][0]) and np.array_equal(h, out[0][1])
        assert np.max(np.abs(w - w_ref)) / np.max(np.abs(w_ref)) < 1e-6
        assert st["batch"] == n and st["mse"] == s_ref["mse"]
    print("THREADS OK %d" % world, flush=True)


if __name__ == "__main__":
    main()
