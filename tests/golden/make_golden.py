#!/usr/bin/env python3
"""Mint the known-answer vectors in tests/golden/*.json.

The reference has NO golden vectors, known-answer tests or fixtures for the learning path
 and cannot run here (JVM/Spark, no JVM in this image), so these are
HAND-DERIVED from the reference's sources and from the published algorithms it calls
(java.lang.String.hashCode, MLlib 1.6.1 HashingTF / LeastSquaresGradient / SimpleUpdater /
GradientDescent.isConverged, Scala BigDecimal HALF_UP).  Every value below is either typed
in literally (with the arithmetic in the comment) or produced by exact rational arithmetic
(fractions.Fraction) and only converted to a double at the end.  Neither oracle/ nor the
product is imported: the oracle is checked AGAINST this file.

Run:  python tests/golden/make_golden.py      (rewrites the JSON next to it)
"""
import json
import math
import os
from fractions import Fraction as Fr

HERE = os.path.dirname(os.path.abspath(__file__))


def dump(name, obj):
    with open(os.path.join(HERE, name), "w") as f:
        json.dump(obj, f, indent=1, sort_keys=True)
        f.write("\n")


# ------------------------------------------------------------------------------------------
# 1. Hashing / featuriser (MllibHelper.scala:42-56,73-82; HashingTF(1000)).
#    String.hashCode of a 2-unit string = 31*c0 + c1; of a 1-unit string = c0.
# ------------------------------------------------------------------------------------------
hash_kats = [
    # units,            java hashCode,  bucket mod 1000
    dict(units=[0x61, 0x62], hash=31 * 97 + 98, bucket=105),          # "ab".hashCode == 3105 (well known)
    dict(units=[0x61], hash=97, bucket=97),                            # "a"
    dict(units=[0x7A, 0x7A], hash=31 * 122 + 122, bucket=904),         # "zz" -> 3904
    dict(units=[0x68, 0xE9], hash=3457, bucket=457),                   # "hé": 31*104+233
    dict(units=[0xD83D, 0xDE00], hash=31 * 0xD83D + 0xDE00, bucket=(31 * 0xD83D + 0xDE00) % 1000),  # the two halves of U+1F600
    dict(units=[0xFFFF, 0xFFFF], hash=32 * 0xFFFF, bucket=(32 * 0xFFFF) % 1000),   # max: 2 097 120, still >= 0
]
assert hash_kats[4]["bucket"] == 899 and hash_kats[5]["hash"] == 2097120

# "héllo 😀!" lower-cased already: h é l l o ' ' D83D DE00 !   (9 UTF-16 units, 8 bigrams)
hello_units = [0x68, 0xE9, 0x6C, 0x6C, 0x6F, 0x20, 0xD83D, 0xDE00, 0x21]
hello_buckets = [457, 331, 456, 459, 473, 349, 899, 825]      # SURVEY.md 8c, re-derived: 31*c0+c1 mod 1000
assert [(31 * a + b) % 1000 for a, b in zip(hello_units, hello_units[1:])] == hello_buckets

NOW = 1464739200000  # 2016-06-01T00:00:00Z


def tweet(units, rc, fol, fav, fri, age_ms, is_rt=1):
    return dict(units=units, retweet_count=rc, followers=fol, favourites=fav, friends=fri,
                created_ms=NOW - age_ms, is_retweet=is_rt)


tweets = [
    tweet(hello_units, 100, 12345, 0, 77, 3600_000),            # kept (lower bound inclusive)
    tweet([0x61, 0x62, 0x61, 0x62, 0x61], 1000, 1, 2, 3, 1),    # "ababa": ab x2, ba x2 ; kept (upper bound inclusive)
    tweet([0x61], 500, 0, 0, 0, 0),                             # 1-unit text -> one 1-unit term; zeros stored explicitly
    tweet([], 500, 5, 5, 5, 5),                                 # empty text -> only the 4 numeric slots
    tweet([0x61, 0x62], 99, 1, 1, 1, 1),                        # dropped: below range
    tweet([0x61, 0x62], 1001, 1, 1, 1, 1),                      # dropped: above range
    tweet([0x61, 0x62], 500, 1, 1, 1, 1, is_rt=0),              # dropped: not a retweet
    tweet([0x7A] * 140, 777, 2147483647, 2147483647, 2147483647, 7 * 86400_000),  # 139 identical bigrams
]

# expected CSR rows, typed by hand
exp_rows = [
    # sorted buckets of hello_buckets, each once
    dict(idx=sorted(hello_buckets) + [1000, 1001, 1002, 1003],
         val=[1.0] * 8 + [12345 * 1e-12, 0 * 1e-12, 77 * 1e-12, 3600_000 * 1e-14], label=100.0),
    # "ab" -> 105, "ba" -> 31*98+97 = 3135 -> 135
    dict(idx=[105, 135, 1000, 1001, 1002, 1003],
         val=[2.0, 2.0, 1 * 1e-12, 2 * 1e-12, 3 * 1e-12, 1 * 1e-14], label=1000.0),
    dict(idx=[97, 1000, 1001, 1002, 1003], val=[1.0, 0.0, 0.0, 0.0, 0.0], label=500.0),
    dict(idx=[1000, 1001, 1002, 1003], val=[5e-12, 5e-12, 5e-12, 5 * 1e-14], label=500.0),
    dict(idx=[904, 1000, 1001, 1002, 1003],
         val=[139.0, 2147483647 * 1e-12, 2147483647 * 1e-12, 2147483647 * 1e-12,
              (7 * 86400_000) * 1e-14], label=777.0),
]
dump("featurizer_kat.json", dict(now_ms=NOW, rt_begin=100, rt_end=1000, num_text_features=1000,
                                 hash_kats=hash_kats, tweets=tweets, expected_rows=exp_rows))

# 1b. Edge cases of the text half, each its own call (own HashingTF width F):
#     a lone high surrogate followed by a letter is just two UTF-16 units (String.sliding works on chars, not code
#     points); a lone low surrogate alone is a one-unit text; the empty text has no term at all; F = 1 sends every
#     term to bucket 0; F = 7 is a small prime width.
edge_groups = [
    dict(num_text_features=1000, tweets=[
        tweet([0xD83D, 0x61], 100, 1, 1, 1, 1),      # 31*55357 + 97 = 1716164 -> 164
        tweet([0xDE00], 100, 2, 2, 2, 2),            # one unit: hashCode = 56832 -> 832
        tweet([], 1000, 3, 3, 3, 3),                 # L = 0: numeric slots only
        tweet([0x61, 0xD83D, 0xDE00, 0xDC00], 500, 4, 4, 4, 4),   # a, a surrogate pair, then a stray low surrogate
    ], expected_rows=[
        dict(idx=[164, 1000, 1001, 1002, 1003], val=[1.0, 1e-12, 1e-12, 1e-12, 1e-14], label=100.0),
        dict(idx=[832, 1000, 1001, 1002, 1003], val=[1.0, 2e-12, 2e-12, 2e-12, 2 * 1e-14], label=100.0),
        dict(idx=[1000, 1001, 1002, 1003], val=[3e-12, 3e-12, 3e-12, 3 * 1e-14], label=1000.0),
        # 31*97+55357 = 58364 -> 364 ; 31*55357+56832 = 1772899 -> 899 ; 31*56832+56320 = 1818112 -> 112
        dict(idx=[112, 364, 899, 1000, 1001, 1002, 1003], val=[1.0, 1.0, 1.0, 4e-12, 4e-12, 4e-12, 4 * 1e-14], label=500.0),
    ]),
    dict(num_text_features=1, tweets=[
        tweet([0x61, 0x62, 0x63], 200, 0, 0, 0, 0),  # "abc": both bigrams in bucket 0
        tweet([0x7A], 300, 0, 0, 0, 0),              # "z": its single one-unit term in bucket 0
        tweet([], 400, 0, 0, 0, 0),
    ], expected_rows=[
        dict(idx=[0, 1, 2, 3, 4], val=[2.0, 0.0, 0.0, 0.0, 0.0], label=200.0),
        dict(idx=[0, 1, 2, 3, 4], val=[1.0, 0.0, 0.0, 0.0, 0.0], label=300.0),
        dict(idx=[1, 2, 3, 4], val=[0.0, 0.0, 0.0, 0.0], label=400.0),
    ]),
    dict(num_text_features=7, tweets=[
        tweet([0x61, 0x62, 0x63], 200, 7, 0, 0, 0),  # ab = 3105 = 7*443 + 4 ; bc = 3137 = 7*448 + 1
    ], expected_rows=[
        dict(idx=[1, 4, 7, 8, 9, 10], val=[1.0, 1.0, 7e-12, 0.0, 0.0, 0.0], label=200.0),
    ]),
]
assert (31 * 0xD83D + 0x61) % 1000 == 164 and 0xDE00 % 1000 == 832
assert [(31 * a + b) % 1000 for a, b in zip([0x61, 0xD83D, 0xDE00], [0xD83D, 0xDE00, 0xDC00])] == [364, 899, 112]
dump("featurizer_edge_kat.json", dict(now_ms=NOW, rt_begin=100, rt_end=1000, groups=edge_groups))

# ------------------------------------------------------------------------------------------
# 2. Utils.round (Utils.scala:4-5): decimal HALF_UP, ties away from zero.
# ------------------------------------------------------------------------------------------
round_kats = [
    [0.5, 1.0], [-0.5, -1.0], [1.5, 2.0], [-1.5, -2.0], [2.5, 3.0], [-2.5, -3.0],
    [0.49999999999999994, 0.0], [-0.49999999999999994, 0.0], [0.0, 0.0], [-0.0, 0.0],
    [-0.3, 0.0], [1e15 + 0.5, 1e15 + 1.0], [4503599627370497.0, 4503599627370497.0],
    [123.456, 123.0], [-123.5, -124.0], [21236.49, 21236.0], [1e300, 1e300],
]
dump("round_kat.json", dict(cases=round_kats))

# ------------------------------------------------------------------------------------------
# 3. runMiniBatchSGD closed forms (LeastSquaresGradient + SimpleUpdater), exact rationals
#    except for 1/sqrt(i), which enters as the double the JVM would compute.
#      diff_r = x_r.w - y_r;  g = sum diff_r x_r;  loss_i = sum diff_r^2 / (2 m)
#      w <- w - (step / sqrt(i)) * g / m
# ------------------------------------------------------------------------------------------


def sgd_exact(X, y, w0, step, iters):
    w = [Fr(v) for v in w0]
    hist, ws = [], []
    m = len(y)
    for i in range(1, iters + 1):
        diffs = [sum(Fr(a) * b for a, b in zip(x, w)) - Fr(t) for x, t in zip(X, y)]
        g = [sum(df * Fr(x[j]) for df, x in zip(diffs, X)) for j in range(len(w))]
        hist.append(float(sum(df * df for df in diffs) / (2 * m)))
        eta = Fr(step / math.sqrt(i))            # the double step/sqrt(i), then exact
        w = [wj - eta * gj / m for wj, gj in zip(w, g)]
        ws.append([float(v) for v in w])
    return ws, hist


cases = []
# (a) one row, two features: by hand  w1 = 0.5*3*(1,2) = (1.5, 3), loss1 = 9/2
ws, hist = sgd_exact([[1, 2]], [3], [0, 0], 0.5, 3)
assert ws[0] == [1.5, 3.0] and hist[0] == 4.5 and hist[1] == 10.125
cases.append(dict(name="one_row", X=[[1, 2]], y=[3], w0=[0, 0], step=0.5, iters=3, tol=0.0,
                  weights_per_iter=ws, loss_history=hist))
# (b) two rows, three features, non-zero start
Xb, yb = [[1, 0, 2], [0.5, -1, 0.25]], [2, -1]
ws, hist = sgd_exact(Xb, yb, [0.125, -0.5, 1], 0.25, 4)
cases.append(dict(name="two_rows", X=Xb, y=yb, w0=[0.125, -0.5, 1], step=0.25, iters=4, tol=0.0,
                  weights_per_iter=ws, loss_history=hist))
# (c) convergence: a tiny step makes ||dw|| < 0.001*max(||w||,1) at i = 2 -> exactly 2 iterations
ws, hist = sgd_exact([[1, 1]], [1], [0, 0], 1e-4, 2)
dn = math.hypot(ws[1][0] - ws[0][0], ws[1][1] - ws[0][1])
assert dn < 0.001 * max(math.hypot(*ws[1]), 1.0)
cases.append(dict(name="converges_at_2", X=[[1, 1]], y=[1], w0=[0, 0], step=1e-4, iters=50, tol=0.001,
                  weights_per_iter=ws, loss_history=hist))
# (d) empty batch: weights untouched, zero iterations (runMiniBatchSGD returns early)
cases.append(dict(name="empty", X=[], y=[], w0=[0.5, -0.5], step=0.1, iters=50, tol=0.001,
                  weights_per_iter=[], loss_history=[]))
dump("sgd_kat.json", dict(cases=cases))

# ------------------------------------------------------------------------------------------
# 4. Predict + statistics block (LinearRegression.scala:56-65), by hand.
#    w = (2, -1); rows x -> raw p -> HALF_UP:  (1,0)->2->2 ; (0.75,0)->1.5->2 ; (0,2.5)->-2.5->-3 ;
#    (0.1,0)->0.2->0 ;  y = (3, 2, -1, 1)
#    (y-p)^2 = 1, 0, 4, 1 -> mse 1.5 -> 2 ; y: mean 1.25, pop.var = 2.1875 -> sd 1.479 -> 1 ;
#    p: (2,2,-3,0) mean .25, pop.var = (3.0625*2 + 10.5625 + .0625)/4 = 4.1875 -> sd 2.046 -> 2
# ------------------------------------------------------------------------------------------
stats = dict(w=[2, -1], X=[[1, 0], [0.75, 0], [0, 2.5], [0.1, 0]], y=[3, 2, -1, 1],
             pred_raw=[2.0, 1.5, -2.5, 0.2], pred_rounded=[2.0, 2.0, -3.0, 0.0],
             batch=4, mse_raw=1.5, mse=2.0, real_stdev_raw=math.sqrt(2.1875), real_stdev=1.0,
             pred_stdev_raw=math.sqrt(4.1875), pred_stdev=2.0)
dump("stats_kat.json", stats)
# ------------------------------------------------------------------------------------------
# 5. One interval of the k-means app (KMeans.scala:93-105), by hand, exact rationals.
#    Points (already scaled; no StandardScaler) (0,0) (0,2) (10,0) (10,2) (11,1); centres c0 = (1,1), c1 = (9,1),
#    weights (2, 0); decayFactor 1/2, "batches".
#    Closest: p0, p1 -> c0 (dist 2 vs 82); p2, p3, p4 -> c1 (dist 2, 2, 4 vs 82, 82, 100).
#    Discount: weights -> (1, 0).  c0: count 2, updated weight 3, lambda 2/3, mean (0,1): c0 = (1/3)(1,1) + (2/3)(0,1) = (1/3, 1)
#    c1: count 3, updated weight 3, lambda 1, mean (31/3, 1): c1 = (31/3, 1).  No dying cluster (3 vs 3).
#    Predict with the updated centres: p0, p1 -> 0; p2, p3, p4 -> 1.
#    StandardScaler(false, true) on the raw columns x = (0,0,10,10,11), y = (0,2,0,2,1): unbiased variances
#    x: mean 6.2, sum sq dev = 38.44*2 + 14.44*2 + 23.04 = 128.8 -> var 32.2 ; y: mean 1, sum sq dev 4 -> var 1.
# ------------------------------------------------------------------------------------------
km_pts = [[0, 0], [0, 2], [10, 0], [10, 2], [11, 1]]
lam0 = Fr(2, 3)
c0 = [float((1 - lam0) * 1 + lam0 * 0), float((1 - lam0) * 1 + lam0 * 1)]
kmeans = dict(points=km_pts, centers=[[1, 1], [9, 1]], weights=[2, 0], decay=0.5, time_unit="batches",
              closest_before=[0, 0, 1, 1, 1], centers_after=[c0, [float(Fr(31, 3)), 1.0]], weights_after=[3.0, 3.0],
              closest_after=[0, 0, 1, 1, 1], column_std=[math.sqrt(32.2), 1.0],
              # the same interval with "points": discount = (1/2)^5 -> weights (1/16, 0); c0: updated 33/16, lambda 32/33
              centers_after_points=[[float(1 - Fr(32, 33)), 1.0], [float(Fr(31, 3)), 1.0]], weights_after_points=[2.0625, 3.0],
              # a dying cluster: centres (1,1), (100,100), weights (4, 0): nobody joins c1 -> weights (2 + 5, 0) ->
              # 0 < 1e-8 * 7: split c0 (updated first: lambda 5/7, mean (6.2, 1) -> c0 = (2/7)(1,1) + (5/7)(6.2, 1) = (33/7, 1))
              dying=dict(centers=[[1, 1], [100, 100]], weights=[4, 0], weights_after=[3.5, 3.5],
                         center_mid=[float(Fr(2, 7) + Fr(5, 7) * Fr(31, 5)), 1.0]))
dump("kmeans_kat.json", kmeans)
print("golden vectors written to", HERE)
