"""GPU parity tests, sparse side: the device featuriser (bit-exact CSR against the oracle and the
hand-derived golden rows) and the CSR training / prediction path, through the C ABI."""
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


def make_model(tw, ctx, d, w0, step=0.1, iters=20, tol=0.0, frac=1.0):
    m = tw.StreamingLinearRegressionWithSGD(stepSize=step, numIterations=iters, miniBatchFraction=frac, ctx=ctx)
    m.setConvergenceTol(tol).setInitialWeights(tw.Vectors.dense(w0))
    return m


def tweet_batch(tw, tweets):
    return tw.TweetBatch([np.array(t["units"], np.uint16) for t in tweets], [t["retweet_count"] for t in tweets],
                         [t["followers"] for t in tweets], [t["favourites"] for t in tweets], [t["friends"] for t in tweets],
                         [t["created_ms"] for t in tweets], [t["is_retweet"] for t in tweets])


def test_featurizer_golden_rows(tw, ctx):
    g = golden("featurizer_kat.json")
    b = ctx.batch_from_tweets(tweet_batch(tw, g["tweets"]), g["now_ms"], g["rt_begin"], g["rt_end"], g["num_text_features"])
    exp = g["expected_rows"]
    assert b.count() == len(exp) and b.d == 1004                       # row count: filter is exact
    rowptr, colidx, vals, labels = b.export_csr()
    for r, e in enumerate(exp):
        sl = slice(rowptr[r], rowptr[r + 1])
        assert colidx[sl].tolist() == e["idx"] and vals[sl].tolist() == e["val"] and labels[r] == e["label"]


ALPHABET = ("abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789 #@.,!?:/"
            "àáâãäåçèéêëìíñöü"
            "\U0001F600\U0001F601\U0001F602\U0001F923\U0001F60D\U0001F44D\U0001F525\U0001F389")


def synth_tweets(n, seed, lower=True):
    """SURVEY.md 8d config 1 recipe: text U[2,140] units over the alphabet, retweet ~ Bernoulli(.7),
    log-uniform counts, created within 7 days before a fixed now."""
    rng = np.random.default_rng(seed)
    now = 1464739200000
    alpha = list(ALPHABET)
    texts = []
    for _ in range(n):
        L = int(rng.integers(0, 141))
        s = "".join(rng.choice(alpha, size=L))
        texts.append(s.lower() if lower else s)
    lu = lambda lo, hi: np.floor(np.exp(rng.uniform(np.log(lo), np.log(hi), n))).astype(np.int64)
    return dict(texts=texts, retweet_count=lu(1, 1e5).astype(np.int32), followers=lu(1, 1e8).astype(np.int32),
                favourites=(lu(1, 1e5) - 1).astype(np.int32), friends=(lu(1, 1e5) - 1).astype(np.int32),
                created_ms=now - rng.integers(0, 7 * 86400_000, n), is_retweet=(rng.random(n) < 0.7).astype(np.uint8)), now


@pytest.mark.parametrize("n,F", [(1, 1000), (257, 1000), (5000, 1000), (3000, 1 << 20), (2000, 7)])
def test_featurizer_bit_exact_vs_oracle(tw, ctx, oracle, n, F):
    t, now = synth_tweets(n, seed=20160601 + n)
    tb = tw.TweetBatch(t["texts"], t["retweet_count"], t["followers"], t["favourites"], t["friends"], t["created_ms"], t["is_retweet"])
    b = ctx.batch_from_tweets(tb, now, 100, 1000, F)
    ref = oracle.featurize(tb.text_utf16, tb.text_off, tb.retweet_count, tb.followers, tb.favourites, tb.friends,
                           tb.created_ms, tb.is_retweet, now, 100, 1000, F)
    got = b.export_csr()
    assert b.count() == len(ref[3]) and b.d == F + 4
    for a, r in zip(got, ref):
        assert a.dtype == r.dtype and np.array_equal(a, r)            # every byte of the CSR


def test_featurizer_edges(tw, ctx, oracle):
    now = 1464739200000
    # all rows filtered out; empty input; maximum-length text; over-long text is refused
    tb = tw.TweetBatch(["ab", "cd"], [5, 5000], [1, 1], [1, 1], [1, 1], [now, now], [1, 1])
    b = ctx.batch_from_tweets(tb, now)
    assert b.count() == 0 and b.isEmpty()
    tb0 = tw.TweetBatch([], [], [], [], [], [], [])
    assert ctx.batch_from_tweets(tb0, now).count() == 0
    long_text = "".join(chr(97 + (i * 7) % 26) for i in range(1025))
    tb1 = tw.TweetBatch([long_text], [500], [3], [4], [5], [now - 9], [1])
    got = ctx.batch_from_tweets(tb1, now).export_csr()
    ref = oracle.featurize(tb1.text_utf16, tb1.text_off, tb1.retweet_count, tb1.followers, tb1.favourites, tb1.friends,
                           tb1.created_ms, tb1.is_retweet, now)
    assert all(np.array_equal(a, r) for a, r in zip(got, ref))
    tb2 = tw.TweetBatch([long_text + "x"], [500], [3], [4], [5], [now], [1])
    with pytest.raises(tw.TwtmlError) as e:
        ctx.batch_from_tweets(tb2, now)
    assert e.value.code == tw._lib.E_INVALID_ARG


def test_library_lowercase_flag(tw, ctx):
    now = 1464739200000
    t, _ = synth_tweets(500, seed=3, lower=False)
    mk = lambda texts: tw.TweetBatch(texts, t["retweet_count"], t["followers"], t["favourites"], t["friends"], t["created_ms"], t["is_retweet"])
    a = ctx.batch_from_tweets(mk(t["texts"]), now, lowercase=True).export_csr()
    b = ctx.batch_from_tweets(mk([s.lower() for s in t["texts"]]), now).export_csr()
    assert all(np.array_equal(x, y) for x, y in zip(a, b))


def test_mllib_helper_shadowing_quirk(tw, ctx):
    """MllibHelper.reset (MllibHelper.scala:22-32) never changes the hashing width."""
    tw.MllibHelper.reset(numRetweetBegin=100, numRetweetEnd=1000, numTextFeatures=5000)
    t, now = synth_tweets(300, seed=8)
    tb = tw.TweetBatch(t["texts"], t["retweet_count"], t["followers"], t["favourites"], t["friends"], t["created_ms"], t["is_retweet"])
    b = tw.MllibHelper.featurize_batch(ctx, tb, now)
    assert b.d == 1004 and tw.MllibHelper.numFeatures == 1004
    tw.MllibHelper.reset()


@pytest.mark.parametrize("n,dt", [(1, 50), (700, 2000), (30000, 1000), (20000, 1 << 20)])
def test_csr_train_matches_oracle(tw, ctx, oracle, n, dt):
    rowptr, colidx, vals, labels = oracle.synth_csr(n, dt, seed=3)
    d = dt + 4
    w0 = np.zeros(d)
    m = make_model(tw, ctx, d, w0, step=0.005, iters=10)
    b = ctx.batch_from_csr(n, d, rowptr, colidx, vals, labels)
    m.trainOnBatch(b)
    w_ref, h_ref = oracle.sgd(labels, d, w0, csr=(rowptr, colidx, vals), step_size=0.005, num_iterations=10,
                              convergence_tol=0.0, partitions=3)
    assert len(m.lastLossHistory) == 10 and relerr(m.lastLossHistory, h_ref) < RTOL
    assert relerr(m.latestModel().weights.toArray(), w_ref) < RTOL
    assert relerr(m.latestModel().weights.toArray(), w_ref) < 1e-11
    # the device generator makes the same bytes
    gb = ctx.batch_synth_csr(n, dt, seed=3)
    for a, r in zip(gb.export_csr(), (rowptr, colidx, vals, labels)):
        assert np.array_equal(a, r)


def test_csr_stats_fraction_and_step(tw, ctx, oracle):
    n, dt = 9000, 1000
    rowptr, colidx, vals, labels = oracle.synth_csr(n, dt, seed=11)
    d = dt + 4
    w = np.random.default_rng(0).normal(0, 30, d)
    b = ctx.batch_from_csr(n, d, rowptr, colidx, vals, labels)
    m = make_model(tw, ctx, d, w, step=0.001, iters=6, frac=0.5)
    st, real, pred = m.predictStats(b)
    s_ref, raw_ref, rnd_ref = oracle.predict_stats(labels, d, w, csr=(rowptr, colidx, vals), partitions=5)
    assert relerr(m.latestModel().predict(b), raw_ref) < 1e-12
    assert np.array_equal(pred, rnd_ref) and np.array_equal(real, labels)
    for k in ("batch", "mse", "real_stdev", "pred_stdev"):
        assert st[k] == s_ref[k]
    for k in ("mse_raw", "real_stdev_raw", "pred_stdev_raw"):
        assert abs(st[k] - s_ref[k]) <= RTOL * abs(s_ref[k])
    m2 = make_model(tw, ctx, d, w, step=0.001, iters=6, frac=0.5)
    st2, _, pred2 = m2.step(b)
    m.trainOnBatch(b)
    assert st2 == st and np.array_equal(pred2, pred)
    assert np.array_equal(m.latestModel().weights.toArray(), m2.latestModel().weights.toArray())
    w_ref, h_ref = oracle.sgd(labels, d, w, csr=(rowptr, colidx, vals), step_size=0.001, num_iterations=6,
                              convergence_tol=0.0, mini_batch_fraction=0.5)
    assert relerr(m.lastLossHistory, h_ref) < RTOL and relerr(m.latestModel().weights.toArray(), w_ref) < RTOL


def test_tweets_end_to_end_like_the_reference(tw, ctx, oracle):
    """featurise -> stats with the pre-training weights -> train, 3 intervals, weights carried over,
    reference defaults (LinearRegression.scala:28-32; reference.conf:3-8)."""
    d = 1004
    model = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(50).setStepSize(0.005)
             .setMiniBatchFraction(1.0).setInitialWeights(tw.Vectors.zeros(d)))
    w_ref = np.zeros(d)
    for it in range(3):
        t, now = synth_tweets(4000, seed=100 + it)
        tb = tw.TweetBatch(t["texts"], t["retweet_count"], t["followers"], t["favourites"], t["friends"], t["created_ms"], t["is_retweet"])
        b = tw.MllibHelper.featurize_batch(ctx, tb, now)
        rp, ci, va, la = oracle.featurize(tb.text_utf16, tb.text_off, tb.retweet_count, tb.followers, tb.favourites,
                                          tb.friends, tb.created_ms, tb.is_retweet, now)
        assert b.count() == len(la)
        st, real, pred = model.step(b)
        s_ref, _, rnd_ref = oracle.predict_stats(la, d, w_ref, csr=(rp, ci, va))
        assert np.array_equal(pred, rnd_ref) and np.array_equal(real, la)
        for k in ("batch", "mse", "real_stdev", "pred_stdev"):
            assert st[k] == s_ref[k]
        w_ref, h_ref = oracle.sgd(la, d, w_ref, csr=(rp, ci, va), step_size=0.005, num_iterations=50)
        assert len(model.lastLossHistory) == len(h_ref)
        assert relerr(model.latestModel().weights.toArray(), w_ref) < RTOL


def test_predict_on_values_keeps_keys_and_order(tw, ctx):
    d = 6
    m = make_model(tw, ctx, d, np.arange(1, d + 1, dtype=float))
    stream = tw.DStream()
    out = []
    m.predictOnValues(stream).foreachRDD(out.append)
    pairs = [("a", tw.Vectors.dense([1, 0, 0, 0, 0, 0])), ("b", tw.Vectors.sparse(d, [1, 5], [2.0, 0.5])),
             ("c", tw.Vectors.dense([0, 0, 0, 0, 0, 0]))]
    stream.push(pairs)
    assert out[0] == [("a", 1.0), ("b", 7.0), ("c", 0.0)]
    assert m.latestModel().predict(tw.Vectors.sparse(d, [0, 2], [1.0, 1.0])) == 4.0
    with pytest.raises(ValueError):
        tw.Vectors.sparse(4, [2, 1], [1.0, 1.0])
    with pytest.raises(tw.TwtmlError):
        ctx.batch_from_csr(1, 4, np.array([0, 2]), np.array([3, 3], np.int32), np.ones(2), np.zeros(1))
