"""GPU tests of the asynchronous ingest paths (the step before the hot path: the reference's receiver fills
the RDD of interval t+1 while foreachRDD works on interval t, LinearRegression.scala:44-47,53,86): every
asynchronous form must leave exactly the bytes the blocking form leaves and train to exactly the same
weights, also when the next batch is being ingested while the current one trains."""
import numpy as np
import pytest

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


def make_model(tw, ctx, d, step=0.005, iters=8):
    m = tw.StreamingLinearRegressionWithSGD(stepSize=step, numIterations=iters, ctx=ctx)
    return m.setConvergenceTol(0.0).setInitialWeights(tw.Vectors.zeros(d))


def pinned(a):
    import torch
    raw = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).pin_memory().numpy()
    return raw.view(a.dtype).reshape(a.shape)


def train(tw, ctx, b, d, **kw):
    m = make_model(tw, ctx, d, **kw)
    st, real, pred = m.step(b)
    return st, pred, m.latestModel().weights.toArray(), m.lastLossHistory


@pytest.mark.parametrize("n,d_text,pin", [(5000, 1000, False), (5000, 1000, True), (3000, 30000, False), (0, 1000, False)])
def test_csr_async_equals_blocking(tw, ctx, oracle, n, d_text, pin):
    rowptr, colidx, vals, labels = oracle.synth_csr(n, d_text, seed=21)
    d = d_text + 4
    arrays = [rowptr, colidx, vals, labels]
    if pin:
        arrays = [pinned(a) for a in arrays]
    bs = ctx.batch_from_csr(n, d, *arrays)
    ba = ctx.batch_from_csr_async(n, d, *arrays)
    assert ba.count() == n                                  # known at once
    assert ba.ready(wait=True)
    for x, y in zip(bs.export_csr(), ba.export_csr()):
        assert x.dtype == y.dtype and np.array_equal(x, y)
    if n:
        rs, ra = train(tw, ctx, bs, d), train(tw, ctx, ba, d)
        assert rs[0] == ra[0] and np.array_equal(rs[1], ra[1]) and np.array_equal(rs[2], ra[2]) and np.array_equal(rs[3], ra[3])
    bs.free()
    ba.free()


@pytest.mark.parametrize("n,F", [(4000, 1000), (2500, 1 << 20), (0, 1000)])
def test_tweets_async_equals_blocking(tw, ctx, n, F):
    from test_sparse_gpu import synth_tweets
    t, now = synth_tweets(n, seed=77 + n)
    tb = tw.TweetBatch(t["texts"], t["retweet_count"], t["followers"], t["favourites"], t["friends"], t["created_ms"], t["is_retweet"])
    bs = ctx.batch_from_tweets(tb, now, 100, 1000, F)
    ba = ctx.batch_from_tweets_async(tb, now, 100, 1000, F)
    assert ba.count() == bs.count()                         # the host filter's count, before the featuriser has run
    assert ba.nnz == bs.nnz
    for x, y in zip(bs.export_csr(), ba.export_csr()):
        assert np.array_equal(x, y)
    if bs.count():
        rs, ra = train(tw, ctx, bs, F + 4), train(tw, ctx, ba, F + 4)
        assert rs[0] == ra[0] and np.array_equal(rs[2], ra[2]) and np.array_equal(rs[3], ra[3])
    bs.free()
    ba.free()


@pytest.mark.parametrize("shift", [0, 1, 3, 8])
@pytest.mark.parametrize("window", [(100, 1000), (1, 10 ** 6)])
def test_tweet_text_read_in_place_from_page_locked_memory(tw, ctx, oracle, shift, window):
    """When few tweets pass the filter and the text column is page-locked, the featuriser reads the kept tweets' text where
    it lies (no copy of the column); with many kept tweets the column is copied.  Either way, and at every alignment of the
    column's first unit, the CSR is the restatement's, byte for byte."""
    from test_sparse_gpu import synth_tweets
    t, now = synth_tweets(6000, seed=900 + shift)
    tb = tw.TweetBatch(t["texts"], t["retweet_count"], t["followers"], t["favourites"], t["friends"], t["created_ms"], t["is_retweet"])
    ref = oracle.featurize(tb.text_utf16, tb.text_off, tb.retweet_count, tb.followers, tb.favourites, tb.friends,
                           tb.created_ms, tb.is_retweet, now, window[0], window[1], 1000)
    room = pinned(np.zeros(tb.text_utf16.size + shift + 8, np.uint16))
    text = room[shift:shift + tb.text_utf16.size]
    text[:] = tb.text_utf16
    room[:shift] = 0xFFFF                                    # units next to the column that must never be looked at
    room[shift + tb.text_utf16.size:] = 0xFFFF
    cols = [pinned(a) for a in (tb.text_off, tb.retweet_count, tb.followers, tb.favourites, tb.friends, tb.created_ms, tb.is_retweet)]
    ptb = tw.TweetBatch.from_arrays(text, *cols)
    assert ptb.text_utf16.ctypes.data == text.ctypes.data   # no copy on the way in
    for make in (ctx.batch_from_tweets, ctx.batch_from_tweets_async):
        b = make(ptb, now, window[0], window[1], 1000)
        got = b.export_csr()
        assert b.count() == len(ref[3])
        for a, r in zip(got, ref):
            assert a.dtype == r.dtype and np.array_equal(a, r)
        b.free()


@pytest.mark.parametrize("pin", [True, False])
def test_dense_rows_appended_in_pieces(tw, ctx, oracle, pin):
    n, d = 20011, 1004
    X, y = oracle.synth_dense(n, d, seed=31)
    y = np.round(y * 300 + 400)
    Xh, yh = (pinned(X), pinned(y)) if pin else (X, y)
    whole = ctx.batch_from_dense(X, y)
    parts = ctx.batch_dense_begin(n, d)
    with pytest.raises(tw.TwtmlError):                       # not all rows in yet
        make_model(tw, ctx, d).step(parts)
    for lo, hi in [(0, 7), (7, 9000), (9000, 9000), (9000, n)]:
        parts.append(Xh[lo:hi], yh[lo:hi])
    with pytest.raises(tw.TwtmlError):
        parts.append(Xh[:1], yh[:1])                          # one row too many
    assert parts.ready(wait=True)
    Xa, ya = parts.export_dense()
    assert np.array_equal(Xa, X) and np.array_equal(ya, y)
    rs, ra = train(tw, ctx, whole, d, step=0.1), train(tw, ctx, parts, d, step=0.1)
    assert rs[0] == ra[0] and np.array_equal(rs[1], ra[1]) and np.array_equal(rs[2], ra[2]) and np.array_equal(rs[3], ra[3])
    whole.free()
    parts.free()


def test_next_interval_is_ingested_while_this_one_trains(tw, ctx, oracle):
    """Three intervals: batch t+1 is handed to the asynchronous ingest before batch t is trained on; weights after the
    third interval equal those of the strictly sequential blocking run, bit for bit."""
    from test_sparse_gpu import synth_tweets
    F, d = 1000, 1004
    batches = []
    for t in range(3):
        tt, now = synth_tweets(6000, seed=500 + t)
        batches.append((tw.TweetBatch(tt["texts"], tt["retweet_count"], tt["followers"], tt["favourites"], tt["friends"], tt["created_ms"],
                                      tt["is_retweet"]), now))
    m_seq = make_model(tw, ctx, d, iters=20)
    stats_seq = []
    for tb, now in batches:
        b = ctx.batch_from_tweets(tb, now, 100, 1000, F)
        stats_seq.append(m_seq.step(b)[0])
        b.free()
    m_pipe = make_model(tw, ctx, d, iters=20)
    stats_pipe = []
    nxt = ctx.batch_from_tweets_async(*batches[0], 100, 1000, F)
    for t in range(3):
        cur = nxt
        nxt = ctx.batch_from_tweets_async(*batches[t + 1], 100, 1000, F) if t + 1 < 3 else None
        stats_pipe.append(m_pipe.step(cur)[0])
        cur.free()
    assert stats_pipe == stats_seq
    assert np.array_equal(m_pipe.latestModel().weights.toArray(), m_seq.latestModel().weights.toArray())


def test_async_failure_is_reported_at_first_use(tw, ctx):
    rowptr = np.array([0, 2, 4], np.int64)
    colidx = np.array([0, 1, 3, 2], np.int32)               # second row: indices decrease
    vals = np.ones(4)
    with pytest.raises(tw.TwtmlError):
        ctx.batch_from_csr(2, 5, rowptr, colidx, vals, np.zeros(2))
    b = ctx.batch_from_csr_async(2, 5, rowptr, colidx, vals, np.zeros(2))     # handed back at once
    with pytest.raises(tw.TwtmlError) as e:
        make_model(tw, ctx, 5).step(b)
    assert e.value.code == tw._lib.E_INVALID_ARG and "strictly increasing" in e.value.message
    b.free()
    # a malformed row pointer is refused before any index is read through it
    bad = np.array([0, 10 ** 9, 4], np.int64)
    with pytest.raises(tw.TwtmlError) as e:
        ctx.batch_from_csr(2, 5, bad, colidx, vals, np.zeros(2))
    assert "rowptr" in e.value.message
    # freeing a batch whose ingest may still be running is fine
    tt = np.arange(1, dtype=np.int64)
    b2 = ctx.batch_from_csr_async(0, 5, tt * 0, np.zeros(0, np.int32), np.zeros(0), np.zeros(0))
    b2.free()
