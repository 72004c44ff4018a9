#!/usr/bin/env python3
"""Tiny invocation of every kernel family (dense CP=1/2/4, row lanes > 1, stats, CSR, featuriser, synthetic
generators, static-scratch stats) for compute-sanitizer runs:
    compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import twtml_b200 as tw

rng = np.random.default_rng(0)
with tw.Context(0) as ctx:
    for n, d in [(37, 4), (130, 128), (77, 1004), (21, 1500), (9, 3000)]:
        X = rng.normal(0, 1, (n, d))
        y = np.round(rng.uniform(100, 1000, n))
        m = tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(3).setConvergenceTol(0.0).setInitialWeights(tw.Vectors.zeros(d))
        b = ctx.batch_from_dense(X, y)
        m.step(b)
        m.trainOnBatch(b)
        m.predictStats(b)
        m.latestModel().predict(b)
        if d <= 1004:
            m.predictStatsStatic(b)
        b.free()
    g = ctx.batch_synth_dense(300, 1004)
    g.export_dense()
    c = ctx.batch_synth_csr(500, 300)
    rp, ci, va, la = c.export_csr()
    m = tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(3).setStepSize(0.001).setConvergenceTol(0.0).setInitialWeights(tw.Vectors.zeros(304))
    m.step(c)
    m.trainOnBatch(ctx.batch_from_csr(500, 304, rp, ci, va, la))
    wide = ctx.batch_synth_csr(400, 20000)
    mw = tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(2).setStepSize(0.001).setConvergenceTol(0.0).setInitialWeights(tw.Vectors.zeros(20004))
    mw.step(wide)
    texts = ["hello world " * (i % 7) + "x" * (i % 13) for i in range(200)]
    tb = tw.TweetBatch(texts, rng.integers(50, 1500, 200), rng.integers(0, 10**6, 200), rng.integers(0, 10**4, 200),
                       rng.integers(0, 10**4, 200), np.full(200, 1464739000000), rng.integers(0, 2, 200))
    fb = ctx.batch_from_tweets(tb, 1464739200000, lowercase=True)
    mf = tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(3).setStepSize(0.005).setInitialWeights(tw.Vectors.zeros(1004))
    mf.step(fb)
print("sanitize_small ok")
