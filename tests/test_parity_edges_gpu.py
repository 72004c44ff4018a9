"""GPU parity tests for the places where a rounding tie, a degenerate text or an empty sample decides
the result: Utils.round's known answers run through every CUDA kernel form (dense, single-pass CSR for
narrow models, two-pass CSR for wide ones), batches whose predictions are EXACT half-integers (no
near-tie mask: every product and partial sum is representable, so summation order cannot matter), the
featuriser's edge texts, the convergence test's "previous weights" gate, and the Lightning series built
from GPU output.  All through the C ABI (ctypes), against the oracle / the hand-derived goldens."""
import json
import math

import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tw():
    import twtml_b200
    return twtml_b200


@pytest.fixture(scope="module")
def ctx(tw):
    c = tw.Context(0)
    yield c
    c.close()


def make_model(tw, ctx, w0, step=0.1, iters=20, tol=0.0, frac=1.0):
    m = tw.StreamingLinearRegressionWithSGD(stepSize=step, numIterations=iters, miniBatchFraction=frac, ctx=ctx)
    m.setConvergenceTol(tol).setInitialWeights(tw.Vectors.dense(w0))
    return m


def one_entry_rows(ctx, xs, d, col):
    """CSR batch with one stored entry per row: row r = {col: xs[r]}."""
    n = len(xs)
    return ctx.batch_from_csr(n, d, np.arange(n + 1, dtype=np.int64), np.full(n, col, np.int32), np.asarray(xs, np.float64), np.zeros(n))


# ---------------------------------------------------------------- Utils.round known answers on the GPU
# d = 1 dense; d = 1 and d = 1004 CSR (k_csr_narrow); d = 3000 and 2^20 + 4 CSR (k_csr_rows): w = e_col, x = the KAT input,
# so the prediction is exactly the KAT input and out_pred_rounded must be exactly Utils.round of it (Utils.scala:4-5).
@pytest.mark.parametrize("form,d,col", [("dense", 1, 0), ("dense", 1004, 17), ("csr", 1, 0), ("csr", 1004, 1003), ("csr", 2048, 5),
                                        ("csr", 3000, 2999), ("csr", (1 << 20) + 4, 1 << 19)])
def test_round_kats_through_cuda(tw, ctx, form, d, col):
    cases = golden("round_kat.json")["cases"]
    xs = np.array([c[0] for c in cases])
    want = np.array([c[1] for c in cases])
    w = np.zeros(d)
    w[col] = 1.0
    m = make_model(tw, ctx, w)
    if form == "dense":
        X = np.zeros((len(xs), d))
        X[:, col] = xs
        b = ctx.batch_from_dense(X, np.zeros(len(xs)))
    else:
        b = one_entry_rows(ctx, xs, d, col)
    st, real, pred = m.predictStats(b)
    assert pred.tolist() == want.tolist()
    assert not np.any(np.signbit(pred[want == 0.0]))          # BigDecimal has no -0.0
    assert st["batch"] == len(xs) and st["nonfinite"] == 0
    # the fused statistics of iteration 1 take the other kernel form (TRAIN_STATS)
    st2, _, pred2 = make_model(tw, ctx, w, step=0.0, iters=1).step(b)
    assert pred2.tolist() == want.tolist() and st2["batch"] == st["batch"]      # (1e300 among the inputs: the moments overflow)
    b.free()


# ---------------------------------------------------------------- exact ties, many of them
def tie_weights(rng, d):
    return rng.choice([-0.5, -0.25, 0.25, 0.5], d)


@pytest.mark.parametrize("n,d", [(4000, 1004), (3000, 128), (500, 3000)])
def test_exact_half_integer_predictions_dense(tw, ctx, oracle, n, d):
    """x small integers, w in {+-1/4, +-1/2}: every product is a multiple of 1/4 below 2^20 -- exact in any order, fused
    or not -- so the raw predictions are bit-equal to the oracle's and about a quarter of them are exact .5 ties."""
    rng = np.random.default_rng(d)
    X = rng.integers(-3, 4, (n, d)).astype(np.float64)
    y = rng.integers(100, 1001, n).astype(np.float64)
    w = tie_weights(rng, d)
    m = make_model(tw, ctx, w)
    b = ctx.batch_from_dense(X, y)
    s_ref, raw_ref, pred_ref = oracle.predict_stats(y, d, w, X=X, partitions=3)
    ties = np.abs(raw_ref - np.trunc(raw_ref)) == 0.5
    assert ties.sum() > n // 8
    st, real, pred = m.predictStats(b)
    assert np.array_equal(m.latestModel().predict(b), raw_ref)
    assert np.array_equal(pred, pred_ref) and np.array_equal(real, y)
    for k in ("batch", "mse", "real_stdev", "pred_stdev"):
        assert st[k] == s_ref[k], k
    for k in ("mse_raw", "real_stdev_raw", "pred_stdev_raw", "real_mean", "pred_mean"):
        assert abs(st[k] - s_ref[k]) <= 1e-12 * abs(s_ref[k]), k
    b.free()


@pytest.mark.parametrize("n,d_text", [(6000, 1000), (5000, 2044), (4000, 5000), (3000, 1 << 20)])
def test_exact_half_integer_predictions_csr(tw, ctx, oracle, n, d_text):
    """TF counts (small positive integers) against weights in {+-1/4, +-1/2}; the four numeric slots carry weight 0.
    d_text 1000 / 2044 run k_csr_narrow, the wider ones k_csr_rows."""
    rng = np.random.default_rng(d_text)
    rowptr, colidx, vals, labels = oracle.synth_csr(n, d_text, seed=9)
    d = d_text + 4
    w = tie_weights(rng, d)
    w[d_text:] = 0.0
    m = make_model(tw, ctx, w)
    b = ctx.batch_from_csr(n, d, rowptr, colidx, vals, labels)
    s_ref, raw_ref, pred_ref = oracle.predict_stats(labels, d, w, csr=(rowptr, colidx, vals), partitions=4)
    ties = np.abs(raw_ref - np.trunc(raw_ref)) == 0.5
    assert ties.sum() > n // 8
    st, real, pred = m.predictStats(b)
    assert np.array_equal(m.latestModel().predict(b), raw_ref)
    assert np.array_equal(pred, pred_ref) and np.array_equal(real, labels)
    for k in ("batch", "mse", "real_stdev", "pred_stdev"):
        assert st[k] == s_ref[k], k
    st2, _, pred2 = make_model(tw, ctx, w, step=0.005, iters=2).step(b)      # TRAIN_STATS form of the same kernels
    assert np.array_equal(pred2, pred_ref) and st2["mse"] == s_ref["mse"]
    b.free()


# ---------------------------------------------------------------- featuriser edge texts
def test_featurizer_edge_kats(tw, ctx):
    g = golden("featurizer_edge_kat.json")
    for grp in g["groups"]:
        F = grp["num_text_features"]
        t = grp["tweets"]
        tb = tw.TweetBatch([np.array(x["units"], np.uint16) for x in t], [x["retweet_count"] for x in t], [x["followers"] for x in t],
                           [x["favourites"] for x in t], [x["friends"] for x in t], [x["created_ms"] for x in t], [x["is_retweet"] for x in t])
        b = ctx.batch_from_tweets(tb, g["now_ms"], g["rt_begin"], g["rt_end"], F)
        assert b.count() == len(grp["expected_rows"]) and b.d == F + 4
        rowptr, colidx, vals, labels = b.export_csr()
        for r, e in enumerate(grp["expected_rows"]):
            sl = slice(rowptr[r], rowptr[r + 1])
            assert colidx[sl].tolist() == e["idx"] and vals[sl].tolist() == e["val"] and labels[r] == e["label"]
        b.free()


# ---------------------------------------------------------------- isConverged waits for previousWeights [U]
@pytest.mark.parametrize("form", ["dense", "csr"])
def test_convergence_gate_with_an_empty_first_sample(tw, ctx, oracle, form):
    """A one-row batch whose row is not drawn at iteration 1: the first non-empty iteration (i = 2) has no previous
    weights to compare with, so the loop must go on to i = 3 and stop there (two loss entries, like the oracle)."""
    from test_oracle import first_sample_empty_case
    from oracle import oracle_np as onp
    frac = first_sample_empty_case(onp)
    X, y = np.array([[1.0, 1.0]]), np.array([1.0])
    w_ref, h_ref = oracle.sgd(y, 2, np.zeros(2), X=X, step_size=1e-4, num_iterations=50, mini_batch_fraction=frac, convergence_tol=0.001)
    assert len(h_ref) == 2
    m = make_model(tw, ctx, np.zeros(2), step=1e-4, iters=50, tol=0.001, frac=frac)
    if form == "dense":
        b = ctx.batch_from_dense(X, y)
    else:
        b = ctx.batch_from_csr(1, 2, np.array([0, 2], np.int64), np.array([0, 1], np.int32), np.array([1.0, 1.0]), y)
    m.trainOnBatch(b)
    assert len(m.lastLossHistory) == 2
    assert np.max(np.abs(m.lastLossHistory - h_ref)) <= 1e-12 * np.max(np.abs(h_ref))
    assert np.max(np.abs(m.latestModel().weights.toArray() - w_ref)) <= 1e-12 * np.max(np.abs(w_ref))
    b.free()


# ---------------------------------------------------------------- what the sinks receive (SessionStats.scala:22-33)
def test_gpu_stats_through_the_wire_formats(tw, ctx, oracle):
    n, d = 300, 1004
    rng = np.random.default_rng(5)
    X = rng.normal(0, 1, (n, d))
    y = np.round(rng.uniform(100, 1000, n))
    w = rng.normal(0, 8, d)
    m = make_model(tw, ctx, w, step=0.005, iters=3)
    b = ctx.batch_from_dense(X, y)
    st, real, pred = m.step(b)
    s_ref, _, pred_ref = oracle.predict_stats(y, d, w, X=X)
    series = tw.lightning_series(st, real, pred)
    assert len(series) == 4 and all(len(s) == n for s in series)
    assert np.array_equal(series[0], y) and np.array_equal(series[1], pred_ref)
    assert np.all(series[2] == s_ref["real_stdev"]) and np.all(series[3] == s_ref["pred_stdev"])
    body = json.loads(tw.stats_payload(12345 + n, st))
    assert body == {"jsonClass": "Stats", "count": 12345 + n, "batch": n, "mse": int(s_ref["mse"]),
                    "realStddev": int(s_ref["real_stdev"]), "predStddev": int(s_ref["pred_stdev"])}
    assert all(isinstance(v, int) for k, v in body.items() if k != "jsonClass")
    assert math.isfinite(st["mse_raw"])
    b.free()
