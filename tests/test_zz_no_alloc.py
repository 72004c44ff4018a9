"""GPU test of twtml_predict_stats_static, the entry point for callers that cannot allocate GPU
memory (a host whose own pool owns the device, or one that cannot take a cudaMalloc stall inside a
batch interval).  It works out of one fixed array inside the library that every call reuses, so the
checks are: the same numbers as twtml_predict_stats (which allocates its scratch) AND as the CPU
oracle -- rounded statistics and counts exactly, raw ones to the suite's 1e-6 relative -- on two calls
in a row with different batches and weights (the second call sees the first one's leftovers in the
array); no change in the device's free memory across calls, the property the entry point exists for,
with the static call as the FIRST compute call of a fresh context; and a clear refusal beyond the
stated size limit.  Collected last (file name) so that a -x run has already run every other test; it
must also pass alone in a fresh process."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-6
EXACT = ("batch", "mse", "real_stdev", "pred_stdev")
RAW = ("mse_raw", "real_stdev_raw", "pred_stdev_raw", "real_mean", "pred_mean")


def _stats(L, fn, m, b):
    import twtml_b200 as tw
    st = tw._lib.Stats()
    if fn == "twtml_predict_stats":
        tw._lib.check(L.twtml_predict_stats(m, b, None, None, C.byref(st)))
    else:
        tw._lib.check(L.twtml_predict_stats_static(m, b, C.byref(st)))
    return st.as_dict()


def _same(got, want):
    for k in EXACT:
        assert got[k] == want[k], (k, got[k], want[k])
    for k in RAW:
        assert abs(got[k] - want[k]) <= RTOL * abs(want[k]), (k, got[k], want[k])


def test_static_scratch_stats_match_allocating_entry_point_twice(oracle):
    import twtml_b200 as tw
    L = tw._lib.lib()
    rng = np.random.default_rng(123)
    with tw.Context(0) as ctx:
        # second case is smaller than the first: stale tuples of call 1 stay in the array beyond its blocks
        for n, d, scale in [(40000, 1004, 6.0), (7001, 1004, 40.0), (5000, 128, 3.0)]:
            X = rng.normal(0, 1, (n, d))
            y = np.round(rng.uniform(100, 1000, n))
            w = rng.normal(0, scale, d)
            model = tw.StreamingLinearRegressionWithSGD(ctx=ctx).setInitialWeights(tw.Vectors.dense(w))
            batch = ctx.batch_from_dense(X, y)
            want = _stats(L, "twtml_predict_stats", model._m, batch._h)
            ref, _, _ = oracle.predict_stats(y, d, w, X=X, partitions=4)
            got1 = _stats(L, "twtml_predict_stats_static", model._m, batch._h)
            got2 = _stats(L, "twtml_predict_stats_static", model._m, batch._h)       # two calls in a row
            for got in (got1, got2):
                assert got["nonfinite"] == want["nonfinite"] == 0
                _same(got, want)
                _same(got, ref)
            assert model.predictStatsStatic(batch) == got2
            batch.free()


def test_static_scratch_allocates_nothing(oracle):
    """The static call as the first compute call of a fresh context: after one warm-up call (address look-up, kernel
    loading), calls on batches of growing size leave the device's free memory exactly where it was."""
    import torch

    import twtml_b200 as tw
    L = tw._lib.lib()
    rng = np.random.default_rng(7)
    d = 1004
    free_bytes = lambda: torch.cuda.mem_get_info(0)[0]
    with tw.Context(0) as ctx:
        w = rng.normal(0, 5.0, d)
        model = tw.StreamingLinearRegressionWithSGD(ctx=ctx).setInitialWeights(tw.Vectors.dense(w))
        data = []
        for n in (3000, 30000, 120000):                      # ingest (which does allocate) happens before the measurement
            X = rng.normal(0, 1, (n, d))
            y = np.round(rng.uniform(100, 1000, n))
            data.append((X, y, ctx.batch_from_dense(X, y)))
        _stats(L, "twtml_predict_stats_static", model._m, data[0][2]._h)          # first compute call on this context: warm-up
        torch.cuda.synchronize()
        before = free_bytes()
        for X, y, b in data:
            got = _stats(L, "twtml_predict_stats_static", model._m, b._h)
            ref, _, _ = oracle.predict_stats(y, d, w, X=X, partitions=2)
            _same(got, ref)
            assert free_bytes() == before, "twtml_predict_stats_static changed the device's free memory"
        for _, _, b in data:
            b.free()


def test_static_scratch_refuses_what_it_cannot_hold():
    import twtml_b200 as tw
    L = tw._lib.lib()
    with tw.Context(0) as ctx:
        model = tw.StreamingLinearRegressionWithSGD(ctx=ctx).setInitialWeights(tw.Vectors.zeros(4))
        big = ctx.batch_synth_dense(16385 * 512 + 8 * 512, 4)        # more than 16384 row blocks of 512 rows
        st = tw._lib.Stats()
        rc = L.twtml_predict_stats_static(model._m, big._h, C.byref(st))
        assert rc == tw._lib.E_INVALID_ARG and "static scratch holds 16384" in tw._lib.last_error()
        assert _stats(L, "twtml_predict_stats", model._m, big._h)["batch"] == big.n     # the allocating call still serves it
        big.free()
        small = ctx.batch_synth_csr(100, 50)
        model2 = tw.StreamingLinearRegressionWithSGD(ctx=ctx).setInitialWeights(tw.Vectors.zeros(54))
        assert L.twtml_predict_stats_static(model2._m, small._h, C.byref(st)) == tw._lib.E_INVALID_ARG
        empty = ctx.batch_from_dense(np.zeros((0, 4)), np.zeros(0))
        assert L.twtml_predict_stats_static(model._m, empty._h, C.byref(st)) == 0 and st.batch == 0
