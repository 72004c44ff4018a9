"""GPU tests of the one-process device group (twtml_create_all): the reference is one JVM driving local[N]
workers (LinearRegression.scala:36), so a host gets all GPUs of the box behind one handle.  Needs >= 2 GPUs
(skipped otherwise; the driver's 1-GPU box skips them, tools/r02_multi.sh runs them on the multi-GPU box)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tw():
    import twtml_b200
    return twtml_b200


@pytest.fixture(scope="module")
def ngpu():
    import torch
    g = torch.cuda.device_count()
    if g < 2:
        pytest.skip("needs >= 2 GPUs")
    return 8 if g >= 8 else (4 if g >= 4 else 2)


@pytest.fixture(scope="module")
def group(tw, ngpu):
    c = tw.Context.all_devices(range(ngpu))
    yield c
    c.close()


@pytest.fixture(scope="module")
def single(tw):
    c = tw.Context(0)
    yield c
    c.close()


def model(tw, ctx, d, step=0.1, iters=12, tol=0.0, frac=1.0):
    m = tw.StreamingLinearRegressionWithSGD(stepSize=step, numIterations=iters, miniBatchFraction=frac, ctx=ctx)
    return m.setConvergenceTol(tol).setInitialWeights(tw.Vectors.zeros(d))


def test_group_reports_its_devices(group, ngpu):
    assert group.device_count() == ngpu


@pytest.mark.parametrize("n,d", [(8 * 1250, 1004), (4096, 128), (64, 1004)])
def test_dense_group_equals_single_gpu_bit_for_bit(tw, group, single, oracle, n, d):
    X, y = oracle.synth_dense(n, d, seed=11)
    y = np.round(y * 300 + 400)
    out = []
    for ctx in (single, group):
        m = model(tw, ctx, d)
        b = ctx.batch_from_dense(X, y)
        assert b.count() == n
        st, real, pred = m.step(b)
        raw = m.latestModel().predict(b)
        out.append((st, real, pred, m.latestModel().weights.toArray(), m.lastLossHistory, raw))
        Xe, ye = b.export_dense()
        assert np.array_equal(Xe, X) and np.array_equal(ye, y)
        b.free()
    a, g = out
    assert a[0] == g[0]
    for i in range(1, 6):
        assert np.array_equal(a[i], g[i]), i               # rows back in the caller's order, weights the same bits
    w_ref, h_ref = oracle.sgd(y, d, np.zeros(d), X=X, step_size=0.1, num_iterations=12, convergence_tol=0.0)
    assert np.max(np.abs(g[3] - w_ref)) / np.max(np.abs(w_ref)) < 1e-6


def test_dense_group_async_and_appended(tw, group, oracle):
    import torch
    n, d = 8 * 700, 1004
    X, y = oracle.synth_dense(n, d, seed=12)
    Xp = torch.from_numpy(X).pin_memory().numpy()
    yp = torch.from_numpy(y).pin_memory().numpy()
    ws = []
    for how in ("blocking", "async", "pieces"):
        m = model(tw, group, d, iters=6, frac=0.5)
        if how == "blocking":
            b = group.batch_from_dense(X, y)
        elif how == "async":
            b = group.batch_from_dense_async(Xp, yp)
        else:
            b = group.batch_dense_begin(n, d)
            for lo, hi in [(0, 100), (100, 3000), (3000, n)]:       # pieces that straddle the members' row blocks
                b.append(Xp[lo:hi], yp[lo:hi])
        st, _, _ = m.step(b)
        assert st["batch"] == n
        ws.append(m.latestModel().weights.toArray())
        b.free()
    assert np.array_equal(ws[0], ws[1]) and np.array_equal(ws[0], ws[2])
    w_ref, _ = oracle.sgd(y, d, np.zeros(d), X=X, step_size=0.1, num_iterations=6, mini_batch_fraction=0.5, convergence_tol=0.0)
    assert np.max(np.abs(ws[0] - w_ref)) / np.max(np.abs(w_ref)) < 1e-6


@pytest.mark.parametrize("d_text", [1000, 30000])
def test_csr_group_matches_oracle(tw, group, oracle, d_text):
    n, d = 8 * 600, d_text + 4
    rowptr, colidx, vals, labels = oracle.synth_csr(n, d_text, seed=13)
    ws = []
    for wait in (True, False):
        m = model(tw, group, d, step=0.005, iters=8)
        b = group.batch_from_csr(n, d, rowptr, colidx, vals, labels, wait=wait)
        assert b.count() == n and b.nnz == len(colidx)
        got = b.export_csr()
        for x, r in zip(got, (rowptr, colidx, vals, labels)):
            assert np.array_equal(x, r)
        st, real, pred = m.step(b)
        ws.append((m.latestModel().weights.toArray(), pred, m.lastLossHistory))
        b.free()
    assert np.array_equal(ws[0][0], ws[1][0])
    w_ref, h_ref = oracle.sgd(labels, d, np.zeros(d), csr=(rowptr, colidx, vals), step_size=0.005, num_iterations=8, convergence_tol=0.0)
    s_ref, _, pred_ref = oracle.predict_stats(labels, d, np.zeros(d), csr=(rowptr, colidx, vals))
    assert np.max(np.abs(ws[0][0] - w_ref)) / np.max(np.abs(w_ref)) < 1e-6
    assert np.max(np.abs(ws[0][2] - h_ref) / np.abs(h_ref)) < 1e-6
    assert np.array_equal(ws[0][1], pred_ref)


def test_tweets_group_keeps_row_order(tw, group, single):
    from test_sparse_gpu import synth_tweets
    t, now = synth_tweets(9000, seed=99)
    tb = tw.TweetBatch(t["texts"], t["retweet_count"], t["followers"], t["favourites"], t["friends"], t["created_ms"], t["is_retweet"])
    bs = single.batch_from_tweets(tb, now)
    for wait in (True, False):
        bg = group.batch_from_tweets(tb, now, wait=wait)
        assert bg.count() == bs.count()
        for x, y in zip(bs.export_csr(), bg.export_csr()):
            assert np.array_equal(x, y)                     # the members' row blocks, concatenated, are the single-GPU batch
        ms, mg = model(tw, single, 1004, step=0.005, iters=10), model(tw, group, 1004, step=0.005, iters=10)
        ss, _, ps = ms.step(bs)
        sg, _, pg = mg.step(bg)
        assert np.array_equal(ps, pg) and ss["batch"] == sg["batch"] and ss["mse"] == sg["mse"]
        a, b = ms.latestModel().weights.toArray(), mg.latestModel().weights.toArray()
        assert np.max(np.abs(a - b)) <= 1e-9 * np.max(np.abs(a))
        bg.free()
    bs.free()
