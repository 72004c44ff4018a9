"""GPU parity tests, dense path: the CUDA kernels, called through the C ABI (ctypes over
libtwtml.so), against the CPU oracle on the same seeded inputs.  Tolerances: weights, loss history
and raw statistics 1e-6 relative (the bar BASELINE.json's north_star states for fp64; observed
~1e-13); row counts, rounded predictions and rounded statistics exact."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu

RTOL = 1e-6


@pytest.fixture(scope="module")
def tw():
    import twtml_b200
    return twtml_b200


@pytest.fixture(scope="module")
def ctx(tw):
    c = tw.Context(0)
    yield c
    c.close()


def relerr(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-300)) if a.size else 0.0


def make_model(tw, ctx, d, w0, step=0.1, iters=20, tol=0.0, frac=1.0, reg=0.0, updater=0):
    m = tw.StreamingLinearRegressionWithSGD(stepSize=step, numIterations=iters, miniBatchFraction=frac, ctx=ctx)
    m.setConvergenceTol(tol).setRegParam(reg).setUpdater(updater).setInitialWeights(tw.Vectors.dense(w0))
    return m


@pytest.mark.parametrize("case", golden("sgd_kat.json")["cases"], ids=lambda c: c["name"])
def test_sgd_golden(tw, ctx, case):
    d = len(case["w0"])
    X = np.array(case["X"], float).reshape(len(case["y"]), d)
    m = make_model(tw, ctx, d, case["w0"], step=case["step"], iters=case["iters"], tol=case["tol"])
    b = ctx.batch_from_dense(X, np.array(case["y"], float))
    m.trainOnBatch(b)
    w = m.latestModel().weights.toArray()
    assert len(m.lastLossHistory) == len(case["loss_history"])        # iterations run, incl. early stop / empty batch
    if case["weights_per_iter"]:
        assert relerr(w, case["weights_per_iter"][-1]) < 1e-12
        assert relerr(m.lastLossHistory, case["loss_history"]) < 1e-12
    else:
        assert w.tolist() == case["w0"]


def test_stats_golden(tw, ctx):
    g = golden("stats_kat.json")
    m = make_model(tw, ctx, 2, g["w"])
    b = ctx.batch_from_dense(np.array(g["X"], float), np.array(g["y"], float))
    st, real, pred = m.predictStats(b)
    assert pred.tolist() == g["pred_rounded"] and real.tolist() == [float(v) for v in g["y"]]
    assert st["batch"] == g["batch"] and st["mse"] == g["mse"]
    assert st["real_stdev"] == g["real_stdev"] and st["pred_stdev"] == g["pred_stdev"]
    assert abs(st["mse_raw"] - g["mse_raw"]) < 1e-14 and abs(st["pred_stdev_raw"] - g["pred_stdev_raw"]) < 1e-14
    raw = m.latestModel().predict(b)
    assert np.max(np.abs(raw - np.array(g["pred_raw"]))) < 1e-15


# (n, d): tiny / ragged row counts, the three column-pass shapes (d=4,128,1004 -> CP=1; 1500 -> CP=2;
# 3000 -> CP=4), row lanes > 1 (d=4, 128), odd d (label shares a column pair with the last feature)
SHAPES = [(1, 4), (7, 4), (1000, 4), (513, 127), (4099, 128), (2500, 1004), (20000, 1004), (777, 1500), (300, 3000),
          (100, 1, ), (3, 2), (90, 1022), (100, 1023), (64, 1024), (60, 1025), (50, 2047), (40, 2048), (35, 2049), (30, 4095), (20, 4096)]   # the width limits of each kernel form


@pytest.mark.parametrize("n,d", SHAPES)
def test_train_matches_oracle(tw, ctx, oracle, n, d):
    X, y = oracle.synth_dense(n, d, seed=100 + d)
    w0 = np.linspace(-0.05, 0.05, d)
    m = make_model(tw, ctx, d, w0, step=0.1, iters=20, tol=0.0)
    b = ctx.batch_from_dense(X, y)
    assert b.count() == n
    m.trainOnBatch(b)
    w_ref, h_ref = oracle.sgd(y, d, w0, X=X, step_size=0.1, num_iterations=20, convergence_tol=0.0, partitions=3, threads=4)
    assert len(m.lastLossHistory) == len(h_ref) == 20
    assert relerr(m.lastLossHistory, h_ref) < RTOL
    assert relerr(m.latestModel().weights.toArray(), w_ref) < RTOL
    # the margin actually achieved is far inside the bar
    assert relerr(m.latestModel().weights.toArray(), w_ref) < 1e-11


def test_dense_width_limit_is_reported(tw, ctx):
    with pytest.raises(tw.TwtmlError) as e:
        ctx.batch_from_dense(np.zeros((2, 4097)), np.zeros(2))
    assert e.value.code == tw._lib.E_INVALID_ARG and "CSR" in e.value.message


def test_ingest_roundtrip_and_synth_bits(tw, ctx, oracle):
    n, d = 3001, 1004
    X, y = oracle.synth_dense(n, d, seed=42, row0=12345)
    b = ctx.batch_from_dense(X, y)
    X2, y2 = b.export_dense()
    assert np.array_equal(X, X2) and np.array_equal(y, y2)
    gb = ctx.batch_synth_dense(n, d, seed=42, row0=12345)
    X3, y3 = gb.export_dense()
    assert np.array_equal(X, X3) and np.array_equal(y, y3)          # device generator == host generator, bit for bit
    # strided host rows (ld > d) and page-locked host memory take other upload paths
    Xw = np.zeros((n, d + 5))
    Xw[:, :d] = X
    b2 = ctx.batch_from_dense(Xw[:, :d], y)
    assert np.array_equal(b2.export_dense()[0], X)
    import torch
    Xp = torch.from_numpy(X).pin_memory()
    b3 = ctx.batch_from_dense(Xp.numpy(), y)
    X4, y4 = b3.export_dense()
    assert np.array_equal(X4, X) and np.array_equal(y4, y)


def test_default_tolerance_stops_early_like_oracle(tw, ctx, oracle):
    X, y = oracle.synth_dense(400, 16, seed=1)
    m = make_model(tw, ctx, 16, np.zeros(16), step=1e-5, iters=50, tol=0.001)
    m.trainOnBatch(ctx.batch_from_dense(X, y))
    w_ref, h_ref = oracle.sgd(y, 16, np.zeros(16), X=X, step_size=1e-5, num_iterations=50)
    assert len(m.lastLossHistory) == len(h_ref) == 2
    assert relerr(m.latestModel().weights.toArray(), w_ref) < RTOL


@pytest.mark.parametrize("updater,reg", [(1, 0.05), (2, 0.01)])
def test_other_updaters(tw, ctx, oracle, updater, reg):
    X, y = oracle.synth_dense(1500, 64, seed=9)
    w0 = np.full(64, 0.02)
    m = make_model(tw, ctx, 64, w0, iters=15, reg=reg, updater=updater)
    m.trainOnBatch(ctx.batch_from_dense(X, y))
    w_ref, h_ref = oracle.sgd(y, 64, w0, X=X, num_iterations=15, convergence_tol=0.0, reg_param=reg, updater=updater)
    assert relerr(m.lastLossHistory, h_ref) < RTOL and relerr(m.latestModel().weights.toArray(), w_ref) < RTOL


def test_mini_batch_fraction(tw, ctx, oracle):
    n, d = 5000, 32
    X, y = oracle.synth_dense(n, d, seed=4)
    m = make_model(tw, ctx, d, np.zeros(d), iters=10, frac=0.25)
    m.trainOnBatch(ctx.batch_from_dense(X, y))
    w_ref, h_ref = oracle.sgd(y, d, np.zeros(d), X=X, num_iterations=10, convergence_tol=0.0, mini_batch_fraction=0.25)
    assert len(m.lastLossHistory) == 10
    assert relerr(m.lastLossHistory, h_ref) < RTOL and relerr(m.latestModel().weights.toArray(), w_ref) < RTOL


@pytest.mark.parametrize("n,d", [(5, 4), (3000, 128), (9000, 1004)])
def test_predict_and_stats_match_oracle(tw, ctx, oracle, n, d):
    X, y = oracle.synth_dense(n, d, seed=77)
    y = np.round(y * 400 + 500)                                      # retweet-count-like labels
    w = np.random.default_rng(5).normal(0, 8, d)
    m = make_model(tw, ctx, d, w)
    b = ctx.batch_from_dense(X, y)
    raw = m.latestModel().predict(b)
    st, real, pred = m.predictStats(b)
    s_ref, raw_ref, rnd_ref = oracle.predict_stats(y, d, w, X=X, partitions=4)
    assert relerr(raw, raw_ref) < 1e-12
    near_tie = np.abs(np.abs(raw_ref - np.trunc(raw_ref)) - 0.5) < 1e-9
    assert np.array_equal(pred[~near_tie], rnd_ref[~near_tie]) and np.array_equal(real, y)
    assert st["batch"] == n == s_ref["batch"]
    for k in ("mse", "real_stdev", "pred_stdev"):
        assert st[k] == s_ref[k], k
    for k in ("mse_raw", "real_stdev_raw", "pred_stdev_raw", "real_mean", "pred_mean"):
        assert abs(st[k] - s_ref[k]) <= RTOL * abs(s_ref[k]), k


def test_step_equals_stats_then_train(tw, ctx, oracle):
    n, d = 6000, 1004
    X, y = oracle.synth_dense(n, d, seed=21)
    y = np.round(y * 300 + 400)
    w0 = np.random.default_rng(1).normal(0, 0.5, d)
    b = ctx.batch_from_dense(X, y)
    m1 = make_model(tw, ctx, d, w0, iters=8)
    st1, real1, pred1 = m1.predictStats(b)
    m1.trainOnBatch(b)
    m2 = make_model(tw, ctx, d, w0, iters=8)
    st2, real2, pred2 = m2.step(b)
    assert st1 == st2 and np.array_equal(pred1, pred2) and np.array_equal(real1, real2)
    assert np.array_equal(m1.latestModel().weights.toArray(), m2.latestModel().weights.toArray())   # bit-identical
    assert np.array_equal(m1.lastLossHistory, m2.lastLossHistory)


def test_weights_carry_across_batches_and_runs_are_bit_reproducible(tw, ctx, oracle):
    d = 1004
    ws = []
    for rep in range(2):
        m = make_model(tw, ctx, d, np.zeros(d), step=0.005, iters=50, tol=0.001)   # reference defaults
        w_ref = np.zeros(d)
        for t in range(3):
            X, y = oracle.synth_dense(1200 + 37 * t, d, seed=50 + t)
            y = np.round(y * 300 + 400)
            m.trainOnBatch(ctx.batch_from_dense(X, y))
            if rep == 0:
                w_ref, h_ref = oracle.sgd(y, d, w_ref, X=X, step_size=0.005, num_iterations=50, partitions=2)
                assert len(m.lastLossHistory) == len(h_ref)
                assert relerr(m.latestModel().weights.toArray(), w_ref) < RTOL
        ws.append(m.latestModel().weights.toArray())
    assert np.array_equal(ws[0], ws[1])


def test_errors_mirror_the_reference(tw, ctx):
    m = tw.StreamingLinearRegressionWithSGD(ctx=ctx)
    with pytest.raises(ValueError, match="Model must be initialized before starting training"):
        m.trainOn(tw.DStream())
    with pytest.raises(ValueError, match="Model must be initialized before starting prediction"):
        m.latestModel()
    m.setInitialWeights(tw.Vectors.zeros(8))
    with pytest.raises(tw.TwtmlError) as e:
        m.trainOnBatch(ctx.batch_from_dense(np.zeros((3, 9)), np.zeros(3)))
    assert e.value.code == tw._lib.E_SHAPE_MISMATCH
    with pytest.raises(tw.TwtmlError):
        m.setConvergenceTol(2.0)


def test_driver_loop_like_the_reference(tw, ctx, oracle):
    """LinearRegression.scala:28-32,53-66,86 replayed: stats op registered before trainOn, so the
    statistics of batch t use the weights trained on batches < t."""
    d = 1004
    model = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(50).setStepSize(0.005)
             .setMiniBatchFraction(1.0).setInitialWeights(tw.Vectors.zeros(d)))
    stream = tw.DStream()
    seen = []
    stream.foreachRDD(lambda rdd: seen.append(None if rdd.isEmpty() else model.predictStats(rdd)[0]))
    model.trainOn(stream)
    w_ref = np.zeros(d)
    for t in range(3):
        X, y = oracle.synth_dense(800, d, seed=200 + t)
        y = np.round(y * 300 + 400)
        stream.push(ctx.batch_from_dense(X, y))
        s_ref, _, _ = oracle.predict_stats(y, d, w_ref, X=X)
        assert seen[-1]["batch"] == 800
        for k in ("mse", "real_stdev", "pred_stdev"):
            assert seen[-1][k] == s_ref[k]
        w_ref, _ = oracle.sgd(y, d, w_ref, X=X, step_size=0.005, num_iterations=50)
    stream.push(ctx.batch_from_dense(np.zeros((0, d)), np.zeros(0)))           # empty interval: both ops skip
    assert seen[-1] is None
    assert relerr(model.latestModel().weights.toArray(), w_ref) < RTOL


def test_async_ingest_overlaps_and_matches_sync(tw, ctx, oracle):
    """twtml_batch_from_dense_async: page-locked rows are copied on the copy stream; the next batch can be
    on its way while the current one trains; contents and results equal the synchronous path's."""
    import torch
    d = 1004
    host = []
    for t in range(3):
        X, y = oracle.synth_dense(3000 + 17 * t, d, seed=400 + t)
        host.append((torch.from_numpy(X).pin_memory().numpy(), torch.from_numpy(np.round(y * 300 + 400)).pin_memory().numpy()))
    m_sync = make_model(tw, ctx, d, np.zeros(d), step=0.005, iters=10)
    m_async = make_model(tw, ctx, d, np.zeros(d), step=0.005, iters=10)
    for X, y in host:
        m_sync.trainOnBatch(ctx.batch_from_dense(X, y))
    nxt = ctx.batch_from_dense_async(*host[0])
    for t in range(3):
        b = nxt
        nxt = ctx.batch_from_dense_async(*host[t + 1]) if t + 1 < 3 else None
        m_async.trainOnBatch(b)
        assert b.ready()
        X2, y2 = b.export_dense()
        assert np.array_equal(X2, host[t][0]) and np.array_equal(y2, host[t][1])
        b.free()
    assert np.array_equal(m_sync.latestModel().weights.toArray(), m_async.latestModel().weights.toArray())
    # pageable memory quietly takes the synchronous path
    Xp, yp = oracle.synth_dense(100, 8, seed=1)
    b = ctx.batch_from_dense_async(Xp, yp)
    assert b.ready(wait=True) and np.array_equal(b.export_dense()[0], Xp)


def test_degenerate_settings_follow_mllib(tw, ctx, oracle):
    """numIterations = 0 runs nothing; miniBatchFraction <= 0 samples no row, so every iteration sees m = 0:
    no update, no loss entry (runMiniBatchSGD's `if (miniBatchSize > 0)` branch [U])."""
    X, y = oracle.synth_dense(500, 16, seed=2)
    w0 = np.linspace(-1, 1, 16)
    b = ctx.batch_from_dense(X, y)
    for iters, frac in [(0, 1.0), (7, 0.0), (7, -0.5)]:
        m = make_model(tw, ctx, 16, w0, iters=iters, frac=frac)
        st, _, _ = m.step(b)
        assert st["batch"] == 500 and len(m.lastLossHistory) == 0
        assert np.array_equal(m.latestModel().weights.toArray(), w0)
        w_ref, h_ref = oracle.sgd(y, 16, w0, X=X, num_iterations=iters, mini_batch_fraction=frac, convergence_tol=0.0)
        assert len(h_ref) == 0 and np.array_equal(w_ref, w0)


def test_nonfinite_prediction_is_flagged(tw, ctx):
    """Utils.round throws NumberFormatException on NaN / Infinity (Utils.scala:4-5): the C ABI counts such rows and
    the facade raises."""
    X = np.ones((3, 4))
    b = ctx.batch_from_dense(X, np.array([1.0, 2.0, 3.0]))
    m = make_model(tw, ctx, 4, np.array([1e308, 1e308, 1e308, 1e308]))
    with pytest.raises(ArithmeticError, match="NumberFormatException"):
        m.predictStats(b)
    assert m.lastStats["nonfinite"] == 3
