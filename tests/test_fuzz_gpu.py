"""Seeded sweep of awkward shapes through the C ABI against the oracle: ragged row counts around the
stage / block / shard boundaries, every dense kernel form, sparse rows of very different lengths
(including empty rows and columns nobody touches), random hyper-parameters."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-6


def relerr(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-300)) if a.size else 0.0


def test_dense_shape_sweep(oracle):
    import twtml_b200 as tw
    rng = np.random.default_rng(2024)
    dims = [1, 2, 3, 5, 31, 32, 33, 63, 64, 65, 127, 129, 255, 500, 511, 512, 513, 1003, 1005, 1100, 2500, 4000]
    with tw.Context(0) as ctx:
        for d in dims:
            for n in sorted({1, 2, int(rng.integers(3, 70)), int(rng.integers(70, 700)), 8 * int(rng.integers(60, 90)) + int(rng.integers(0, 8))}):
                X = rng.normal(0, 1, (n, d)) / np.sqrt(d)
                y = X @ rng.normal(0, 1, d) + rng.normal(0, 0.1, n)
                w0 = rng.normal(0, 0.1, d)
                iters = int(rng.integers(1, 6))
                step = float(rng.uniform(0.01, 0.5))
                frac = float(rng.choice([1.0, 1.0, 0.6]))
                m = (tw.StreamingLinearRegressionWithSGD(stepSize=step, numIterations=iters, miniBatchFraction=frac, ctx=ctx)
                     .setConvergenceTol(0.0).setInitialWeights(tw.Vectors.dense(w0)))
                b = ctx.batch_from_dense(X, y)
                st, real, pred = m.step(b)
                w_ref, h_ref = oracle.sgd(y, d, w0, X=X, step_size=step, num_iterations=iters, mini_batch_fraction=frac,
                                          convergence_tol=0.0, partitions=2)
                s_ref, raw_ref, rnd_ref = oracle.predict_stats(y, d, w0, X=X)
                tag = (n, d, iters, frac)
                assert len(m.lastLossHistory) == len(h_ref), tag
                assert relerr(m.lastLossHistory, h_ref) < RTOL and relerr(m.latestModel().weights.toArray(), w_ref) < RTOL, tag
                near_tie = np.abs(np.abs(raw_ref - np.trunc(raw_ref)) - 0.5) < 1e-9
                assert np.array_equal(pred[~near_tie], rnd_ref[~near_tie]) and np.array_equal(real, y), tag
                assert st["batch"] == n and abs(st["mse_raw"] - s_ref["mse_raw"]) <= RTOL * max(abs(s_ref["mse_raw"]), 1e-300), tag
                b.free()
                m.free()


def test_sparse_shape_sweep(oracle):
    import twtml_b200 as tw
    rng = np.random.default_rng(77)
    with tw.Context(0) as ctx:
        for d in [1, 7, 100, 1004, 5000, 70000]:
            for n in [1, 2, 33, 257, 1500]:
                lens = rng.choice([0, 0, 1, 2, 5, 40, min(d, 300)], size=n)
                lens = np.minimum(lens, d)
                rowptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
                colidx = np.concatenate([np.sort(rng.choice(d, size=k, replace=False)) for k in lens] + [np.zeros(0, np.int64)]).astype(np.int32)
                vals = rng.integers(1, 4, size=colidx.size).astype(np.float64) * rng.choice([1.0, 1e-6], size=colidx.size)
                labels = np.round(rng.uniform(100, 1000, n))
                w0 = rng.normal(0, 0.5, d)
                iters = int(rng.integers(1, 5))
                m = (tw.StreamingLinearRegressionWithSGD(stepSize=0.01, numIterations=iters, ctx=ctx)
                     .setConvergenceTol(0.0).setInitialWeights(tw.Vectors.dense(w0)))
                b = ctx.batch_from_csr(n, d, rowptr, colidx, vals, labels)
                st, real, pred = m.step(b)
                w_ref, h_ref = oracle.sgd(labels, d, w0, csr=(rowptr, colidx, vals), step_size=0.01, num_iterations=iters, convergence_tol=0.0)
                s_ref, raw_ref, rnd_ref = oracle.predict_stats(labels, d, w0, csr=(rowptr, colidx, vals))
                tag = (n, d, iters, int(colidx.size))
                assert relerr(m.lastLossHistory, h_ref) < RTOL and relerr(m.latestModel().weights.toArray(), w_ref) < RTOL, tag
                near_tie = np.abs(np.abs(raw_ref - np.trunc(raw_ref)) - 0.5) < 1e-9
                assert np.array_equal(pred[~near_tie], rnd_ref[~near_tie]) and st["batch"] == n, tag
                got = b.export_csr()
                assert np.array_equal(got[0], rowptr) and np.array_equal(got[1], colidx) and np.array_equal(got[2], vals), tag
                b.free()
                m.free()
