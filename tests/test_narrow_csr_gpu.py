"""GPU parity tests for the single-pass kernel of narrow CSR models (csr_narrow.cuh, d <= 2048): the shapes that
stress its ring of entry tiles and its row bookkeeping -- rows of the maximum length (three tiles per row), rows
longer than the entries a lane keeps in registers, empty rows, fewer rows than warps, one enormous CTA stretch --
against the oracle, and against the library's own two-pass kernels on the same batch (TWTML_NO_NARROW=1)."""
import os

import numpy as np
import pytest

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


def random_csr(n, d, row_len, seed, scale=1.0):
    """row_len(rng, r) -> entries of row r; indices strictly increasing, values ~ N(0, scale), labels ~ N(0, 10)."""
    rng = np.random.default_rng(seed)
    lens = np.array([min(d, max(0, int(row_len(rng, r)))) for r in range(n)], np.int64)
    rowptr = np.zeros(n + 1, np.int64)
    np.cumsum(lens, out=rowptr[1:])
    colidx = np.zeros(int(rowptr[-1]), np.int32)
    for r in range(n):
        colidx[rowptr[r]:rowptr[r + 1]] = np.sort(rng.choice(d, size=int(lens[r]), replace=False))
    vals = rng.normal(0, scale, int(rowptr[-1]))
    labels = rng.normal(0, 10, n)
    return rowptr, colidx, vals, labels


def train(tw, ctx, d, csr, w0, step, iters, frac=1.0, narrow=True):
    rowptr, colidx, vals, labels = csr
    if narrow:
        os.environ.pop("TWTML_NO_NARROW", None)
    else:
        os.environ["TWTML_NO_NARROW"] = "1"
    try:
        b = ctx.batch_from_csr(len(labels), d, rowptr, colidx, vals, labels)
    finally:
        os.environ.pop("TWTML_NO_NARROW", None)
    m = tw.StreamingLinearRegressionWithSGD(stepSize=step, numIterations=iters, miniBatchFraction=frac, ctx=ctx)
    m.setConvergenceTol(0.0).setInitialWeights(tw.Vectors.dense(w0))
    st, real, pred = m.step(b)
    raw = m.latestModel().predict(b)
    b.free()
    return m.latestModel().weights.toArray(), np.array(m.lastLossHistory), st, pred, raw


SHAPES = {
    # name: (n, d, row length, step)
    "max_rows_d2048": (300, 2048, lambda rng, r: 2048, 1e-4),                  # every row spans three tiles
    "long_and_empty_d2048": (500, 2048, lambda rng, r: (0, 2048, 5, 1500)[r % 4], 1e-4),
    "tail_loop_d300": (4000, 300, lambda rng, r: rng.integers(90, 301), 1e-3),  # > 96 entries: beyond the lanes' registers
    "short_rows_d16": (20000, 16, lambda rng, r: rng.integers(0, 9), 1e-2),
    "fewer_rows_than_warps": (5, 1004, lambda rng, r: rng.integers(1, 120), 1e-3),
    "one_row": (1, 2048, lambda rng, r: 1000, 1e-4),
    "all_empty": (64, 40, lambda rng, r: 0, 1e-2),
    "tweet_like_d1004": (60000, 1004, lambda rng, r: rng.integers(5, 144), 5e-4),
    "d1": (1000, 1, lambda rng, r: rng.integers(0, 2), 1e-2),
    # more ring / tail shapes at d <= 1024
    "max_rows_d1024": (301, 1024, lambda rng, r: 1024, 2e-4),            # rows of a whole tile; odd row count
    "long_and_empty_d1024": (999, 1024, lambda rng, r: (0, 1024, 3, 700, 0, 0, 81)[r % 7], 2e-4),
    "tail_loop_d200": (5001, 200, lambda rng, r: rng.integers(70, 201), 1e-3),   # rows around and beyond the 96 entries the lanes keep in registers
    "three_rows": (3, 1004, lambda rng, r: (100, 0, 17)[r], 1e-3),
}


@pytest.mark.parametrize("name", list(SHAPES))
def test_narrow_kernel_matches_oracle_and_two_pass(tw, ctx, oracle, name):
    n, d, row_len, step = SHAPES[name]
    csr = random_csr(n, d, row_len, seed=len(name) * 7919 + n)
    w0 = np.random.default_rng(1).normal(0, 0.1, d)
    w, hist, st, pred, raw = train(tw, ctx, d, csr, w0, step, 8)
    w_ref, h_ref = oracle.sgd(csr[3], d, w0, csr=csr[:3], step_size=step, num_iterations=8, convergence_tol=0.0, partitions=3)
    assert len(hist) == 8 and relerr(hist, h_ref) < RTOL
    assert relerr(w, w_ref) < RTOL
    s_ref, raw_ref, rnd_ref = oracle.predict_stats(csr[3], d, w0, csr=csr[:3], partitions=2)
    assert np.array_equal(pred, rnd_ref)                                       # statistics ride on the first iteration
    for k in ("batch", "mse", "real_stdev", "pred_stdev"):
        assert st[k] == s_ref[k]
    _, raw_w, _ = oracle.predict_stats(csr[3], d, w_ref, csr=csr[:3])
    assert relerr(raw, raw_w) < 1e-6
    if d > 1:  # the library's other CSR path on the same batch
        w2, hist2, st2, pred2, raw2 = train(tw, ctx, d, csr, w0, step, 8, narrow=False)
        assert relerr(w, w2) < 1e-10 and relerr(hist, hist2) < 1e-10
        assert np.array_equal(pred, pred2) and all(st[k] == st2[k] for k in ("batch", "mse", "real_stdev", "pred_stdev"))


def test_narrow_kernel_mini_batch_sampling(tw, ctx, oracle):
    n, d = 30000, 1004
    csr = random_csr(n, d, lambda rng, r: rng.integers(0, 130), seed=5)
    w0 = np.zeros(d)
    w, hist, _, _, _ = train(tw, ctx, d, csr, w0, 1e-3, 6, frac=0.3)
    w_ref, h_ref = oracle.sgd(csr[3], d, w0, csr=csr[:3], step_size=1e-3, num_iterations=6, convergence_tol=0.0,
                              mini_batch_fraction=0.3)
    assert relerr(hist, h_ref) < RTOL and relerr(w, w_ref) < RTOL


def test_narrow_kernel_is_reproducible(tw, ctx):
    n, d = 50000, 700
    csr = random_csr(n, d, lambda rng, r: rng.integers(0, 200), seed=9)
    w0 = np.zeros(d)
    a = train(tw, ctx, d, csr, w0, 1e-3, 5)
    b = train(tw, ctx, d, csr, w0, 1e-3, 5)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])          # bit for bit, run to run
