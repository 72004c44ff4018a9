"""Independent numpy/pure-Python twin of oracle/twtml_oracle.c -- TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED (the reference has no test or fixture on this path and cannot run here; see
the header of twtml_oracle.c).  This twin is written differently on purpose -- decimal
arithmetic for Utils.round, Python str/UTF-16 handling and a Counter for the featuriser,
matrix products for the SGD sums, two-pass moments for the statistics -- so that agreement
between the two restatements (tests/test_oracle.py, <= 1e-12 relative) catches a slip in
either.  Citations: file:line under /root/reference/spark/src/main/scala/com/giorgioinf/
twtml/spark/; [U] = Apache Spark MLlib 1.6.1 (un-vendored dependency, spark/build.sbt:15).
"""
import math
from collections import Counter
from decimal import ROUND_HALF_UP, Decimal

import numpy as np


def round_half_up(x):
    """Utils.round (Utils.scala:4-5): BigDecimal(Double.toString(x)).setScale(0, HALF_UP)."""
    x = float(x)
    if not math.isfinite(x):
        return x
    return float(Decimal(repr(x)).to_integral_value(rounding=ROUND_HALF_UP)) + 0.0


def utf16_units(s):
    """Java chars of a Python str (surrogate pairs split, as String.sliding sees them)."""
    b = s.encode("utf-16-le", "surrogatepass")
    return [b[i] | (b[i + 1] << 8) for i in range(0, len(b), 2)]


def java_hash(units):
    h = 0
    for c in units:
        h = (31 * h + c) & 0xFFFFFFFF
    return h - (1 << 32) if h >= (1 << 31) else h


def hash_term(units, num_features):
    """HashingTF.indexOf [U]: nonNegativeMod(term.##, numFeatures)."""
    h = java_hash(units)
    r = int(math.fmod(h, num_features))          # Java %: sign of the dividend
    return r + num_features if r < 0 else r


def featurize_text(lowered, num_text_features=1000):
    """MllibHelper.featurizeText (MllibHelper.scala:42-56) after toLowerCase."""
    u = utf16_units(lowered)
    if len(u) == 0:
        terms = []
    elif len(u) == 1:
        terms = [u]                               # sliding(2) on a 1-char string: one short window
    else:
        terms = [u[i:i + 2] for i in range(len(u) - 1)]
    tf = Counter(hash_term(t, num_text_features) for t in terms)
    idx = sorted(tf)
    return idx, [float(tf[i]) for i in idx]


def featurize(tweets, now_ms, rt_begin=100, rt_end=1000, num_text_features=1000):
    """tweets: iterable of dicts(text, retweet_count, followers, favourites, friends,
    created_ms, is_retweet) with text already lower-cased.  MllibHelper.filtrate (:89-94)
    then featurize (:73-82).  -> rowptr, colidx, vals, labels."""
    rowptr, colidx, vals, labels = [0], [], [], []
    F = num_text_features
    for t in tweets:
        if not (t["is_retweet"] and rt_begin <= t["retweet_count"] <= rt_end):
            continue
        idx, v = featurize_text(t["text"], F)
        nums = [t["followers"] * 1e-12, t["favourites"] * 1e-12, t["friends"] * 1e-12,
                float(now_ms - t["created_ms"]) * 1e-14]          # :58-71
        colidx += idx + [F, F + 1, F + 2, F + 3]
        vals += v + nums
        labels.append(float(t["retweet_count"]))
        rowptr.append(len(colidx))
    return (np.array(rowptr, np.int64), np.array(colidx, np.int32), np.array(vals, np.float64),
            np.array(labels, np.float64))


_M64 = (1 << 64) - 1


def splitmix64(z):
    z = (z + 0x9E3779B97F4A7C15) & _M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def _splitmix64_np(z):
    z = z.astype(np.uint64)
    with np.errstate(over="ignore"):
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def u01_np(seed, idx):
    z = _splitmix64_np(np.uint64(seed) ^ idx.astype(np.uint64))
    return (z >> np.uint64(11)).astype(np.float64) * 2.0 ** -53


def sample_mask(iteration, row0, n, frac):
    """Stand-in for RDD.sample(false, frac, 42+i) [U]; see twtml_oracle.c."""
    if frac >= 1.0:
        return np.ones(n, bool)
    if frac <= 0.0:
        return np.zeros(n, bool)
    seed = splitmix64(42 + iteration)
    return u01_np(seed, np.arange(row0, row0 + n, dtype=np.uint64)) < frac


def synth_dense(n, d, seed=42, row0=0):
    inv = 1.0 / math.sqrt(d)
    r = np.arange(row0, row0 + n, dtype=np.uint64)[:, None]
    j = np.arange(d, dtype=np.uint64)[None, :]
    with np.errstate(over="ignore"):
        X = (2.0 * u01_np(seed, r * np.uint64(d) + j) - 1.0) * inv
    ws = (2.0 * u01_np(seed + 1, np.arange(d, dtype=np.uint64)) - 1.0) * inv
    acc = np.zeros(n)
    for c in range(d):                            # index order, unfused, like the C loop
        acc = acc + X[:, c] * ws[c]
    y = acc + 0.01 * (2.0 * u01_np(seed + 2, r[:, 0]) - 1.0)
    return X, y


def _densify(csr, n, d):
    rowptr, colidx, vals = csr
    X = np.zeros((n, d))
    for r in range(n):
        X[r, colidx[rowptr[r]:rowptr[r + 1]]] = vals[rowptr[r]:rowptr[r + 1]]
    return X


def _update(kind, w, g, step, it, reg):
    """Updaters [U] optimization/Updater.scala -> (w', regVal)."""
    ts = step / math.sqrt(it)
    if kind == 0:
        return w - ts * g, 0.0
    if kind == 1:
        w2 = w * (1.0 - ts * reg) - ts * g
        return w2, 0.5 * reg * float(np.linalg.norm(w2)) ** 2
    w2 = w - ts * g
    w2 = np.sign(w2) * np.maximum(0.0, np.abs(w2) - reg * ts)
    return w2, float(np.abs(w2).sum()) * reg


def sgd(y, d, w0, X=None, csr=None, step_size=0.1, num_iterations=50, mini_batch_fraction=1.0,
        convergence_tol=0.001, reg_param=0.0, updater=0, row0=0):
    """GradientDescent.runMiniBatchSGD [U] with LeastSquaresGradient; one micro-batch."""
    y = np.asarray(y, np.float64)
    n = len(y)
    w = np.array(w0, np.float64)
    hist = []
    if n == 0:
        return w, np.array(hist)
    if X is None:
        X = _densify(csr, n, d)
    _, reg_val = _update(updater, w, np.zeros(d), 0.0, 1, reg_param)
    converged, i = False, 1
    while not converged and i <= num_iterations:
        m = sample_mask(i, row0, n, mini_batch_fraction)
        cnt = int(m.sum())
        if cnt > 0:
            Xs, ys = X[m], y[m]
            diff = Xs @ w - ys
            g = Xs.T @ diff
            hist.append(float((diff * diff).sum() / 2.0) / cnt + reg_val)
            w_new, reg_val = _update(updater, w, g / cnt, step_size, i, reg_param)
            if len(hist) >= 2:        # previousWeights != None [U]: an earlier iteration also drew a non-empty sample
                converged = np.linalg.norm(w - w_new) < convergence_tol * max(np.linalg.norm(w_new), 1.0)
            w = w_new
        i += 1
    return w, np.array(hist)


def predict_stats(y, d, w, X=None, csr=None):
    """LinearRegression.scala:53-66 (predict with pre-training weights, round, stats)."""
    y = np.asarray(y, np.float64)
    n = len(y)
    if X is None:
        X = _densify(csr, n, d)
    raw = X @ np.asarray(w, np.float64) + 0.0
    prnd = np.array([round_half_up(v) for v in raw])
    if n == 0:
        nan = float("nan")
        return dict(batch=0, mse=nan, real_stdev=nan, pred_stdev=nan), raw, prnd
    mse = float(np.mean((y - prnd) ** 2))
    rs, ps = float(np.std(y)), float(np.std(prnd))      # population, like StatCounter.stdev [U]
    st = dict(batch=n, mse=round_half_up(mse), real_stdev=round_half_up(rs),
              pred_stdev=round_half_up(ps), mse_raw=mse, real_stdev_raw=rs, pred_stdev_raw=ps,
              real_mean=float(y.mean()), pred_mean=float(prnd.mean()))
    return st, raw, prnd


# ---- KMeans.scala:93-105, written independently of the C file (vectorised; np.var for the scaler) ------------------------
def column_std(X):
    X = np.asarray(X, np.float64)
    return np.sqrt(np.var(X, axis=0, ddof=1)) if X.shape[0] > 1 else np.zeros(X.shape[1])


def scale(X, std):
    inv = np.array([1.0 / s if s != 0.0 else 0.0 for s in std])
    return np.asarray(X, np.float64) * inv


def kmeans_predict(centers, X):
    """Closest centre by squared Euclidean distance, ties to the lower index (exact differences: agrees with MLlib's
    pruned / expanded form away from near-ties)."""
    centers, X = np.asarray(centers, np.float64), np.asarray(X, np.float64)
    d2 = ((X[:, None, :] - centers[None, :, :]) ** 2).sum(axis=2)
    return np.argmin(d2, axis=1).astype(np.int32)


def kmeans_update(centers, weights, X, decay, time_unit=0):
    c, w = np.array(centers, np.float64), np.array(weights, np.float64)
    X = np.asarray(X, np.float64)
    lab = kmeans_predict(c, X)
    w = w * (decay ** len(X) if time_unit == 1 else decay)
    for k in range(len(c)):
        pts = X[lab == k]
        if len(pts) == 0:
            continue
        new_w = w[k] + len(pts)
        lam = len(pts) / max(new_w, 1e-16)
        c[k] = (1.0 - lam) * c[k] + lam * pts.mean(axis=0)
        w[k] = new_w
    big, small = int(np.argmax(w)), int(np.argmin(w))
    if w[small] < 1e-8 * w[big]:
        w[big] = w[small] = (w[big] + w[small]) / 2.0
        p = 1e-14 * np.maximum(np.abs(c[big]), 1.0)
        c[small] = c[big] - p
        c[big] = c[big] + p
    return c, w
