#!/usr/bin/env python3
"""Stress of the narrow-CSR kernel's ring protocol: random shapes for a fixed wall time; every batch is trained twice with the
single-pass kernel (must agree bit for bit) and once with the two-pass kernels (must agree to 1e-10)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import twtml_b200 as tw

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)


def batch(n, d, mean_len):
    lens = np.minimum(d, rng.poisson(mean_len, n) * (rng.random(n) < 0.9)).astype(np.int64)
    if rng.random() < 0.3:
        lens[rng.integers(0, n, max(1, n // 50))] = d  # some rows of the maximum length
    rowptr = np.zeros(n + 1, np.int64)
    np.cumsum(lens, out=rowptr[1:])
    nnz = int(rowptr[-1])
    # strictly increasing indices per row: sorted sample without replacement, vectorised by sorting random keys per row
    colidx = np.empty(nnz, np.int32)
    for r in np.nonzero(lens)[0]:
        colidx[rowptr[r]:rowptr[r + 1]] = np.sort(rng.choice(d, size=int(lens[r]), replace=False))
    return rowptr, colidx, rng.normal(0, 1, nnz), rng.normal(0, 10, n)


def run(ctx, d, csr, narrow, iters=6):
    if not narrow:
        os.environ["TWTML_NO_NARROW"] = "1"
    try:
        b = ctx.batch_from_csr(len(csr[3]), d, *csr)
    finally:
        os.environ.pop("TWTML_NO_NARROW", None)
    m = tw.StreamingLinearRegressionWithSGD(stepSize=1e-4, numIterations=iters, ctx=ctx).setConvergenceTol(0.0)
    m.setInitialWeights(tw.Vectors.zeros(d))
    st, _, pred = m.step(b)
    b.free()
    return m.latestModel().weights.toArray(), np.array(m.lastLossHistory), pred


t0, k = time.time(), 0
with tw.Context(0) as ctx:
    while time.time() - t0 < budget:
        d = int(rng.choice([1, 3, 17, 64, 300, 1004, 1500, 2048]))
        n = int(rng.choice([1, 7, 100, 2000, 20000, 60000]))
        mean_len = float(rng.choice([0.5, 4, 30, 75, 200, 1200]))
        if n * min(d, mean_len) > 6e6:
            continue
        csr = batch(n, d, mean_len)
        a = run(ctx, d, csr, True)
        b = run(ctx, d, csr, True)
        c = run(ctx, d, csr, False)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2]), (n, d, mean_len)
        sc = max(1e-300, float(np.max(np.abs(c[0]))))
        assert np.max(np.abs(a[0] - c[0])) / sc < 1e-10 and np.array_equal(a[2], c[2]), (n, d, mean_len)
        k += 1
print("narrow-CSR stress: %d random batches in %.0f s, all reproducible and equal to the two-pass result" % (k, time.time() - t0))
