#!/usr/bin/env python3
"""Multi-GPU parity check, run under torchrun (one rank per GPU):
   the same micro-batch, row-sharded over G GPUs, must give the SAME BITS as on one GPU when
   n % 8 == 0 (fixed virtual-shard reduction order), must match the CPU oracle to 1e-6, and the
   statistics must be those of the whole batch on every rank.  Prints one line 'MGPU OK ...'."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import twtml_b200 as tw
from twtml_b200.dist import ranked_context, row_range
from oracle import oracle_c


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    ctx = ranked_context(local)
    results = {}
    for n, d, iters in [(8 * 1000, 1004, 12), (4096, 128, 10), (8 * 333 + 3, 1004, 5)]:
        X, y = oracle_c.synth_dense(n, d, seed=5)
        y = np.round(y * 300 + 400)
        w0 = np.linspace(-0.1, 0.1, d)
        lo, hi = row_range(n, rank, world)
        m = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(iters).setStepSize(0.1)
             .setConvergenceTol(0.0).setInitialWeights(tw.Vectors.dense(w0)))
        b = ctx.batch_from_dense(X[lo:hi], y[lo:hi])
        st, real, pred = m.step(b)
        w = m.latestModel().weights.toArray()
        w_ref, h_ref = oracle_c.sgd(y, d, w0, X=X, step_size=0.1, num_iterations=iters, convergence_tol=0.0)
        s_ref, _, rnd_ref = oracle_c.predict_stats(y, d, w0, X=X)
        assert np.max(np.abs(w - w_ref)) / np.max(np.abs(w_ref)) < 1e-6
        assert np.max(np.abs(m.lastLossHistory - h_ref) / np.abs(h_ref)) < 1e-6
        assert st["batch"] == n and st["mse"] == s_ref["mse"] and st["real_stdev"] == s_ref["real_stdev"] and st["pred_stdev"] == s_ref["pred_stdev"]
        assert np.array_equal(pred, rnd_ref[lo:hi]) and np.array_equal(real, y[lo:hi])
        # every rank holds the same weights, bit for bit
        wt = torch.from_numpy(w).cuda()
        w0t = wt.clone()
        dist.broadcast(w0t, src=0)
        assert torch.equal(wt, w0t), "ranks disagree on the weights"
        results[(n, d)] = w
        # sparse form of a small batch too (one shard per rank)
        if d == 128:
            rp, ci, va, la = oracle_c.synth_csr(n, 500, seed=9)
            ms = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(6).setStepSize(0.005)
                  .setConvergenceTol(0.0).setInitialWeights(tw.Vectors.zeros(504)))
            rps = rp[lo:hi + 1] - rp[lo]
            bs = ctx.batch_from_csr(hi - lo, 504, rps, ci[rp[lo]:rp[hi]], va[rp[lo]:rp[hi]], la[lo:hi])
            ms.trainOnBatch(bs)
            ws_ref, hs_ref = oracle_c.sgd(la, 504, np.zeros(504), csr=(rp, ci, va), step_size=0.005, num_iterations=6, convergence_tol=0.0)
            assert np.max(np.abs(ms.latestModel().weights.toArray() - ws_ref)) / np.max(np.abs(ws_ref)) < 1e-6
            wsn = ms.latestModel().weights.toArray()
            wtn = torch.from_numpy(wsn).cuda(); w0n = wtn.clone(); dist.broadcast(w0n, src=0)
            assert torch.equal(wtn, w0n), "ranks disagree on the narrow CSR weights"
            # the same model when only rank 0 has rows (an interval with fewer tweets than GPUs): the other ranks feed empty
            # batches and still take part in every exchange
            k = 3
            me = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(4).setStepSize(0.005)
                  .setConvergenceTol(0.0).setInitialWeights(tw.Vectors.zeros(504)))
            if rank == 0:
                be = ctx.batch_from_csr(k, 504, rp[:k + 1], ci[:rp[k]], va[:rp[k]], la[:k])
            else:
                be = ctx.batch_from_csr(0, 504, np.zeros(1, np.int64), np.zeros(0, np.int32), np.zeros(0), np.zeros(0))
            ste, _, _ = me.step(be)
            we_ref, _ = oracle_c.sgd(la[:k], 504, np.zeros(504), csr=(rp[:k + 1], ci[:rp[k]], va[:rp[k]]), step_size=0.005,
                                     num_iterations=4, convergence_tol=0.0)
            assert ste["batch"] == k
            assert np.max(np.abs(me.latestModel().weights.toArray() - we_ref)) / np.max(np.abs(we_ref)) < 1e-6
            # a wide hashed model takes the sliced exchange (all-to-all of gradient slices, sliced update, all-gather of w)
            dw = 20000 + 4
            rp, ci, va, la = oracle_c.synth_csr(n, 20000, seed=10)
            mw = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(6).setStepSize(0.005)
                  .setInitialWeights(tw.Vectors.zeros(dw)))                      # default tolerance: may stop early
            bw = ctx.batch_from_csr(hi - lo, dw, rp[lo:hi + 1] - rp[lo], ci[rp[lo]:rp[hi]], va[rp[lo]:rp[hi]], la[lo:hi])
            stw, _, _ = mw.step(bw)
            ww_ref, hw_ref = oracle_c.sgd(la, dw, np.zeros(dw), csr=(rp, ci, va), step_size=0.005, num_iterations=6)
            ww = mw.latestModel().weights.toArray()
            assert len(mw.lastLossHistory) == len(hw_ref) and stw["batch"] == n
            assert np.max(np.abs(ww - ww_ref)) / np.max(np.abs(ww_ref)) < 1e-6
            assert np.max(np.abs(mw.lastLossHistory - hw_ref) / np.abs(hw_ref)) < 1e-6
            wt2 = torch.from_numpy(ww).cuda(); w0t2 = wt2.clone(); dist.broadcast(w0t2, src=0)
            assert torch.equal(wt2, w0t2), "ranks disagree on the wide weights"
    # ---- the exchange buffer's two parities across calls: train twice on one resident batch with tolerances that make
    # the first call stop after an odd, then after an even number of exchanges; no collective sits between the calls, so a
    # sequence number that counted enqueued (rather than executed) exchanges would let a fast rank overwrite a parity a
    # slow one is still reading.  Weights must equal the oracle's and be the same bits on every rank, every round.
    n, d = 8 * 64, 16
    X, y = oracle_c.synth_dense(n, d, seed=3)
    lo, hi = row_range(n, rank, world)
    b = ctx.batch_from_dense(X[lo:hi], y[lo:hi])
    for tol, step in [(0.5, 0.1), (0.02, 0.1), (0.5, 0.3), (0.1, 0.05), (0.0, 0.1)]:
        w_ref = np.zeros(d)
        m = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(9).setStepSize(step)
             .setConvergenceTol(tol).setInitialWeights(tw.Vectors.zeros(d)))
        ran = []
        for call in range(4):
            m.trainOnBatch(b)
            w_ref, h_ref = oracle_c.sgd(y, d, w_ref, X=X, step_size=step, num_iterations=9, convergence_tol=tol)
            ran.append(len(h_ref))
            assert len(m.lastLossHistory) == len(h_ref), (tol, call, len(m.lastLossHistory), len(h_ref))
            w = m.latestModel().weights.toArray()
            assert np.max(np.abs(w - w_ref)) <= 1e-9 * max(1.0, np.max(np.abs(w_ref)))
            wt = torch.from_numpy(w).cuda(); w0t = wt.clone(); dist.broadcast(w0t, src=0)
            assert torch.equal(wt, w0t), "ranks disagree after repeated calls"
        results[("rep%g_%g" % (tol, step), 0)] = np.array(ran, np.float64)
    b.free()
    # ---- asynchronous ingest on a ranked context: page-locked rows, the row-count exchange deferred to first use
    n, d = 8 * 500, 1004
    X, y = oracle_c.synth_dense(n, d, seed=8)
    lo, hi = row_range(n, rank, world)
    Xp = torch.from_numpy(np.ascontiguousarray(X[lo:hi])).pin_memory().numpy()
    yp = torch.from_numpy(np.ascontiguousarray(y[lo:hi])).pin_memory().numpy()
    ws = []
    for make in (lambda: ctx.batch_from_dense(Xp, yp), lambda: ctx.batch_from_dense_async(Xp, yp)):
        m = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(6).setStepSize(0.1).setMiniBatchFraction(0.5)
             .setConvergenceTol(0.0).setInitialWeights(tw.Vectors.zeros(d)))
        bb = make()
        st, _, _ = m.step(bb)
        assert st["batch"] == n
        ws.append(m.latestModel().weights.toArray())
        bb.free()
    assert np.array_equal(ws[0], ws[1]), "asynchronous ingest changed the result"
    results[("async", 0)] = ws[1]              # n % 8 == 0: the same bits on any number of GPUs
    w_ref, _ = oracle_c.sgd(y, d, np.zeros(d), X=X, step_size=0.1, num_iterations=6, mini_batch_fraction=0.5, convergence_tol=0.0)
    assert np.max(np.abs(ws[1] - w_ref)) / np.max(np.abs(w_ref)) < 1e-6        # the sampler saw the right global row numbers
    rp, ci, va, la = oracle_c.synth_csr(n, 1000, seed=12)
    parts = (rp[lo:hi + 1] - rp[lo], ci[rp[lo]:rp[hi]], va[rp[lo]:rp[hi]], la[lo:hi])
    ws = []
    for wait in (True, False):
        m = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(5).setStepSize(0.005)
             .setConvergenceTol(0.0).setInitialWeights(tw.Vectors.zeros(1004)))
        bb = ctx.batch_from_csr(hi - lo, 1004, *parts, wait=wait)
        m.trainOnBatch(bb)
        ws.append(m.latestModel().weights.toArray())
        bb.free()
    assert np.array_equal(ws[0], ws[1]), "asynchronous CSR ingest changed the result"
    results[("csr_async", 0)] = ws[1]          # CSR: one shard per rank, equal across GPU counts to rounding only
    if rank == 0:
        np.savez(os.environ.get("MGPU_OUT", "/tmp/mgpu_w%d.npz" % world), **{"%s_%s" % k: v for k, v in results.items()})
        print("MGPU OK world=%d" % world, flush=True)
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
