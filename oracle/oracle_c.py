"""ctypes front for oracle/twtml_oracle.c -- TEST INFRASTRUCTURE ONLY (parity unpinned, see
the C file's header).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libtwtml_oracle.so")
_lib = None

_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("twtml_oracle.c", "twtml_oracle_kmeans.c", "Makefile")]
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(p) for p in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"], env=dict(os.environ))
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.twtml_oracle_round.restype = C.c_double
        _lib.twtml_oracle_round.argtypes = [C.c_double]
        _lib.twtml_oracle_hash_term.restype = C.c_int32
        _lib.twtml_oracle_sampled.restype = C.c_int32
        _lib.twtml_oracle_sampled.argtypes = [C.c_int32, C.c_uint64, C.c_double]
        _lib.twtml_oracle_featurize.restype = C.c_int64
        _lib.twtml_oracle_synth_csr.restype = C.c_int64
        _lib.twtml_oracle_sgd.restype = C.c_int32
        _lib.twtml_oracle_predict_stats.restype = C.c_int32
        _lib.twtml_oracle_max_threads.restype = C.c_int32
    return _lib


def max_threads():
    return int(lib().twtml_oracle_max_threads())


def round_half_up(x):
    return float(lib().twtml_oracle_round(float(x)))


def hash_term(units, num_features):
    u = np.ascontiguousarray(units, dtype=np.uint16)
    return int(lib().twtml_oracle_hash_term(u.ctypes.data_as(C.c_void_p), C.c_int64(len(u)),
                                            C.c_int32(num_features)))


def sampled(iteration, global_row, frac):
    return bool(lib().twtml_oracle_sampled(int(iteration), int(global_row), float(frac)))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def featurize(text_utf16, text_off, retweet_cnt, followers, favourites, friends, created_ms,
              is_retweet, now_ms, rt_begin=100, rt_end=1000, num_text_features=1000):
    """tweets -> (rowptr, colidx, vals, labels) of the kept rows (CSR, width F+4)."""
    L = lib()
    n = len(retweet_cnt)
    text_utf16 = np.ascontiguousarray(text_utf16, np.uint16)
    if text_utf16.size == 0:
        text_utf16 = np.zeros(1, np.uint16)
    args = [np.ascontiguousarray(text_off, np.int64), np.ascontiguousarray(retweet_cnt, np.int32),
            np.ascontiguousarray(followers, np.int32), np.ascontiguousarray(favourites, np.int32),
            np.ascontiguousarray(friends, np.int32), np.ascontiguousarray(created_ms, np.int64),
            np.ascontiguousarray(is_retweet, np.uint8)]
    nnz = C.c_int64(0)

    def call(rowptr, colidx, vals, labels):
        return L.twtml_oracle_featurize(
            _p(text_utf16), *[_p(a) for a in args], C.c_int64(n), C.c_int64(now_ms),
            C.c_int32(rt_begin), C.c_int32(rt_end), C.c_int32(num_text_features),
            _p(rowptr), _p(colidx), _p(vals), _p(labels), C.byref(nnz))

    kept = call(None, None, None, None)
    rowptr = np.zeros(kept + 1, np.int64)
    colidx = np.zeros(max(nnz.value, 1), np.int32)
    vals = np.zeros(max(nnz.value, 1), np.float64)
    labels = np.zeros(max(kept, 1), np.float64)
    call(rowptr, colidx, vals, labels)
    return rowptr, colidx[:nnz.value], vals[:nnz.value], labels[:kept]


def synth_dense(n, d, seed=42, row0=0, threads=None, want_x=True):
    threads = threads or max_threads()
    X = np.empty((n, d), np.float64) if want_x else None
    y = np.empty(n, np.float64)
    lib().twtml_oracle_synth_dense(C.c_int64(n), C.c_int64(d), C.c_uint64(seed), C.c_int64(row0),
                                   _p(X), C.c_int64(d), _p(y), C.c_int(threads))
    return X, y


def synth_csr(n, d_text, seed=42, row0=0):
    rowptr = np.zeros(n + 1, np.int64)
    L = lib()
    nnz = L.twtml_oracle_synth_csr(C.c_int64(n), C.c_int32(d_text), C.c_uint64(seed),
                                   C.c_int64(row0), _p(rowptr), None, None, None)
    colidx = np.zeros(max(nnz, 1), np.int32)
    vals = np.zeros(max(nnz, 1), np.float64)
    labels = np.zeros(max(n, 1), np.float64)
    L.twtml_oracle_synth_csr(C.c_int64(n), C.c_int32(d_text), C.c_uint64(seed), C.c_int64(row0),
                             _p(rowptr), _p(colidx), _p(vals), _p(labels))
    return rowptr, colidx[:nnz], vals[:nnz], labels[:n]


def sgd(y, d, w0, X=None, csr=None, step_size=0.1, num_iterations=50, mini_batch_fraction=1.0,
        convergence_tol=0.001, reg_param=0.0, updater=0, partitions=1, threads=1, row0=0):
    """runMiniBatchSGD on one micro-batch -> (weights, loss_history[iters_run])."""
    y = np.ascontiguousarray(y, np.float64)
    n = len(y)
    w = np.array(w0, np.float64).copy()
    assert w.shape == (d,)
    hist = np.zeros(max(num_iterations, 1), np.float64)
    it = C.c_int32(0)
    if X is not None:
        X = np.ascontiguousarray(X, np.float64)
        ldx, rp, ci, va = (X.shape[1] if X.ndim == 2 else d), None, None, None
    else:
        rp, ci, va = (np.ascontiguousarray(csr[0], np.int64), np.ascontiguousarray(csr[1], np.int32),
                      np.ascontiguousarray(csr[2], np.float64))
        ldx = 0
    rc = lib().twtml_oracle_sgd(
        _p(X), C.c_int64(ldx), _p(rp), _p(ci), _p(va), _p(y), C.c_int64(n), C.c_int64(d), _p(w),
        C.c_double(step_size), C.c_int32(num_iterations), C.c_double(mini_batch_fraction),
        C.c_double(convergence_tol), C.c_double(reg_param), C.c_int32(updater),
        C.c_int32(partitions), C.c_int32(threads), C.c_int64(row0), _p(hist), C.byref(it))
    if rc != 0:
        raise ValueError("twtml_oracle_sgd: bad arguments")
    return w, hist[:it.value].copy()


def sgd_synth(n, d, w0, seed=42, row0=0, step_size=0.1, num_iterations=3, convergence_tol=0.0, threads=None):
    """runMiniBatchSGD over the synthetic dense batch (seed, row0), rows regenerated on the fly."""
    threads = threads or max_threads()
    w = np.array(w0, np.float64).copy()
    hist = np.zeros(max(num_iterations, 1), np.float64)
    it = C.c_int32(0)
    L = lib()
    L.twtml_oracle_sgd_synth.restype = C.c_int32
    rc = L.twtml_oracle_sgd_synth(C.c_int64(n), C.c_int64(d), C.c_uint64(seed), C.c_int64(row0), _p(w),
                                  C.c_double(step_size), C.c_int32(num_iterations), C.c_double(convergence_tol),
                                  C.c_int32(threads * 4), C.c_int32(threads), _p(hist), C.byref(it))
    if rc != 0:
        raise ValueError("twtml_oracle_sgd_synth: bad arguments")
    return w, hist[:it.value].copy()


def predict_stats(y, d, w, X=None, csr=None, partitions=1):
    """-> dict(batch, mse, real_stdev, pred_stdev, *_raw, means), pred_raw[n], pred_rounded[n]."""
    y = np.ascontiguousarray(y, np.float64)
    n = len(y)
    w = np.ascontiguousarray(w, np.float64)
    praw = np.zeros(max(n, 1), np.float64)
    prnd = np.zeros(max(n, 1), np.float64)
    out = np.zeros(9, np.float64)
    if X is not None:
        X = np.ascontiguousarray(X, np.float64)
        ldx, rp, ci, va = (X.shape[1] if X.ndim == 2 else d), None, None, None
    else:
        rp, ci, va = (np.ascontiguousarray(csr[0], np.int64), np.ascontiguousarray(csr[1], np.int32),
                      np.ascontiguousarray(csr[2], np.float64))
        ldx = 0
    rc = lib().twtml_oracle_predict_stats(_p(X), C.c_int64(ldx), _p(rp), _p(ci), _p(va), _p(y),
                                          C.c_int64(n), C.c_int64(d), _p(w), C.c_int32(partitions),
                                          _p(praw), _p(prnd), _p(out))
    if rc != 0:
        raise ValueError("twtml_oracle_predict_stats: bad arguments")
    keys = ["batch", "mse", "real_stdev", "pred_stdev", "mse_raw", "real_stdev_raw",
            "pred_stdev_raw", "real_mean", "pred_mean"]
    st = dict(zip(keys, out.tolist()))
    st["batch"] = int(st["batch"])
    return st, praw[:n], prnd[:n]


# ---- the reference's second app (KMeans.scala:93-105): StandardScaler + StreamingKMeansModel.update / predict -----------
def column_std(X):
    """StandardScaler(false, true).fit(rdd).std [U]: unbiased column standard deviations."""
    X = np.ascontiguousarray(X, np.float64)
    out = np.zeros(X.shape[1])
    lib().twtml_oracle_column_std(_p(X), C.c_int64(X.shape[0]), C.c_int64(X.shape[1]), C.c_int32(X.shape[1]), _p(out))
    return out


def scale(X, std):
    X = np.ascontiguousarray(X, np.float64)
    std = np.ascontiguousarray(std, np.float64)
    out = np.zeros_like(X)
    lib().twtml_oracle_scale(_p(X), C.c_int64(X.shape[0]), C.c_int64(X.shape[1]), C.c_int32(X.shape[1]), _p(std), _p(out))
    return out


def kmeans_predict(centers, X):
    centers = np.ascontiguousarray(centers, np.float64)
    X = np.ascontiguousarray(X, np.float64)
    out = np.zeros(max(X.shape[0], 1), np.int32)
    lib().twtml_oracle_kmeans_predict(_p(centers), C.c_int32(centers.shape[0]), C.c_int32(centers.shape[1]), _p(X),
                                      C.c_int64(X.shape[0]), C.c_int64(X.shape[1]), _p(out))
    return out[:X.shape[0]]


def kmeans_update(centers, weights, X, decay, time_unit=0):
    """StreamingKMeansModel.update [U] -> (new centers, new weights); time_unit 0 batches, 1 points."""
    c = np.array(centers, np.float64).copy()
    w = np.array(weights, np.float64).copy()
    X = np.ascontiguousarray(X, np.float64)
    lib().twtml_oracle_kmeans_update(_p(c), _p(w), C.c_int32(c.shape[0]), C.c_int32(c.shape[1]), _p(X), C.c_int64(X.shape[0]),
                                     C.c_int64(X.shape[1]), C.c_double(decay), C.c_int32(time_unit))
    return c, w
