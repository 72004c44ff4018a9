#!/usr/bin/env python3
"""Latency of one batch interval at the reference's observed operating point (doc/graph.png: ~11 retweets per 5 s
batch): featurise + stats + 50 SGD iterations (reference defaults, convergenceTol 0.001) through the C ABI."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import twtml_b200 as tw

with tw.Context(0) as ctx:
    model = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(50).setStepSize(0.005)
             .setMiniBatchFraction(1.0).setInitialWeights(tw.Vectors.zeros(1004)))
    for n_tweets in (80, 800, 8000, 80000):
        cols, now = bench.synth_tweets(n_tweets, 7)
        tb = tw.TweetBatch.from_arrays(**cols)
        ts = []
        for rep in range(30):
            t0 = time.perf_counter()
            b = ctx.batch_from_tweets(tb, now)
            st, _, _ = model.step(b)
            b.free()
            ts.append(time.perf_counter() - t0)
        ts = sorted(ts[5:])
        print("tweets %6d -> rows %5d: median %.3f ms, p90 %.3f ms per batch interval (iterations run: %d)" % (
            n_tweets, st["batch"], 1e3 * ts[len(ts) // 2], 1e3 * ts[int(len(ts) * 0.9)], len(model.lastLossHistory)), flush=True)
