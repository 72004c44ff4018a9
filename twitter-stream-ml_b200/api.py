"""Host-side mirror of the Scala API the reference drives (LinearRegression.scala:28-32,53-66,86).

The reference's host language is Scala on a JVM, which this image does not have, so the facade a
Spark job would use (JNA binding + Scala class) is mirrored here in Python over the
same C ABI, name for name: StreamingLinearRegressionWithSGD with setStepSize / setNumIterations /
setMiniBatchFraction / setInitialWeights / setConvergenceTol, latestModel(), trainOn, predictOn,
predictOnValues; LinearRegressionModel.predict; Vectors / LabeledPoint; MllibHelper.filtrate /
featurize; and a DStream stand-in that runs registered output operations per micro-batch in
registration order, as Spark Streaming's JobScheduler does.  All arithmetic happens in libtwtml.so on
the GPU; nothing here computes a dot product, a gradient or a statistic.
"""
import ctypes as C
import weakref

import numpy as np

from . import _lib
from ._lib import TwtmlError, check, ptr


# ------------------------------------------------------------------------------------------------
# context, batches
# ------------------------------------------------------------------------------------------------
class Context:
    """One GPU (twtml_ctx).  Context.ranked() makes one rank of a row-sharded multi-GPU job."""

    def __init__(self, device=0, rank=0, world_size=1, nccl_id=None):
        self._h = C.c_void_p()
        L = _lib.lib()
        if isinstance(device, (list, tuple)):               # twtml_create_all: one handle over several GPUs of this process
            devs = (C.c_int32 * len(device))(*device)
            check(L.twtml_create_all(devs, len(device), C.byref(self._h)))
            device = device[0]
        elif world_size == 1:
            check(L.twtml_create(device, C.byref(self._h)))
        else:
            buf = (C.c_uint8 * _lib.NCCL_ID_BYTES).from_buffer_copy(bytes(nccl_id))
            check(L.twtml_create_ranked(device, rank, world_size, buf, _lib.NCCL_ID_BYTES, C.byref(self._h)))
        self.device, self.rank, self.world_size = device, rank, world_size
        self._children = weakref.WeakSet()     # live Batch / model wrappers: freed before the context goes

    @classmethod
    def all_devices(cls, devices):
        """One context driving the listed GPUs from this process (twtml_create_all)."""
        return cls(list(devices))

    def device_count(self):
        n = C.c_int32()
        check(_lib.lib().twtml_device_count(self._h, C.byref(n)))
        return n.value

    @staticmethod
    def nccl_unique_id():
        buf = (C.c_uint8 * _lib.NCCL_ID_BYTES)()
        check(_lib.lib().twtml_nccl_unique_id(buf, _lib.NCCL_ID_BYTES))
        return bytes(buf)

    def profile(self, on=True):
        check(_lib.lib().twtml_profile_enable(self._h, int(on)))

    def last_profile(self):
        p = _lib.Profile()
        check(_lib.lib().twtml_profile_get(self._h, C.byref(p)))
        return p

    def close(self):
        if self._h:
            for ch in list(self._children):
                ch.free()
            check(_lib.lib().twtml_destroy(self._h))
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- micro-batch ingest
    def batch_from_dense(self, X, labels):
        X = np.asarray(X, np.float64)
        labels = np.ascontiguousarray(labels, np.float64)
        if X.ndim != 2 or X.shape[0] != labels.shape[0]:
            raise ValueError("X must be n x d with one label per row")
        if X.strides[1] != 8 or X.strides[0] % 8:
            X = np.ascontiguousarray(X)
        h = C.c_void_p()
        check(_lib.lib().twtml_batch_from_dense(self._h, X.shape[0], X.shape[1], ptr(X), X.strides[0] // 8 if X.shape[0] > 1 else X.shape[1],
                                                ptr(labels), C.byref(h)))
        return Batch(self, h)

    def batch_from_dense_async(self, X, labels):
        """Like batch_from_dense, but returns while the copy runs (page-locked X / labels, C-contiguous
        rows); the arrays are kept alive by the returned Batch until it is ready."""
        X = np.asarray(X, np.float64)
        labels = np.ascontiguousarray(labels, np.float64)
        if X.ndim != 2 or X.shape[0] != labels.shape[0] or X.strides[1] != 8 or X.strides[0] % 8:
            raise ValueError("X must be n x d float64 with unit column stride and one label per row")
        h = C.c_void_p()
        check(_lib.lib().twtml_batch_from_dense_async(self._h, X.shape[0], X.shape[1], ptr(X), X.strides[0] // 8 if X.shape[0] > 1 else X.shape[1],
                                                      ptr(labels), C.byref(h)))
        b = Batch(self, h)
        b._keep = (X, labels)
        return b

    def batch_dense_begin(self, n, d):
        """A dense batch of n rows whose rows arrive with Batch.append(X, labels), piece by piece (the blocks a
        receiver stores during one batch interval); usable once all n rows are in."""
        h = C.c_void_p()
        check(_lib.lib().twtml_batch_dense_begin(self._h, n, d, C.byref(h)))
        b = Batch(self, h)
        b._keep = []
        return b

    def batch_from_csr(self, n, d, rowptr, colidx, vals, labels, wait=True):
        """wait=False: twtml_batch_from_csr_async -- returns at once, the ingest runs beside the context's compute."""
        rowptr, colidx = _lib.arr(rowptr, np.int64), _lib.arr(colidx, np.int32)
        vals, labels = _lib.arr(vals, np.float64), _lib.arr(labels, np.float64)
        h = C.c_void_p()
        fn = _lib.lib().twtml_batch_from_csr if wait else _lib.lib().twtml_batch_from_csr_async
        check(fn(self._h, n, d, ptr(rowptr), ptr(colidx), ptr(vals), ptr(labels), C.byref(h)))
        b = Batch(self, h)
        if not wait:
            b._keep = (rowptr, colidx, vals, labels)
        return b

    def batch_from_csr_async(self, n, d, rowptr, colidx, vals, labels):
        return self.batch_from_csr(n, d, rowptr, colidx, vals, labels, wait=False)

    def batch_from_tweets(self, tweets, now_ms, rt_begin=100, rt_end=1000, num_text_features=1000, lowercase=False, wait=True):
        """tweets: TweetBatch (struct of arrays).  Filter + featurise on the GPU.  wait=False:
        twtml_batch_from_tweets_async (the row count is known at once, nnz once the featuriser has run)."""
        t = tweets
        h = C.c_void_p()
        text = t.text_utf16 if t.text_utf16.size else np.zeros(1, np.uint16)
        fn = _lib.lib().twtml_batch_from_tweets if wait else _lib.lib().twtml_batch_from_tweets_async
        check(fn(self._h, ptr(text), ptr(t.text_off), ptr(t.retweet_count), ptr(t.followers), ptr(t.favourites),
                 ptr(t.friends), ptr(t.created_ms), ptr(t.is_retweet), len(t), int(now_ms), rt_begin, rt_end,
                 num_text_features, _lib.FLAG_LOWERCASE if lowercase else 0, C.byref(h)))
        b = Batch(self, h)
        if not wait:
            b._keep = (text, t)
        return b

    def batch_from_tweets_async(self, tweets, now_ms, rt_begin=100, rt_end=1000, num_text_features=1000, lowercase=False):
        return self.batch_from_tweets(tweets, now_ms, rt_begin, rt_end, num_text_features, lowercase, wait=False)

    def batch_synth_dense(self, n, d, seed=42, row0=0):
        h = C.c_void_p()
        check(_lib.lib().twtml_batch_synth_dense(self._h, n, d, seed, row0, C.byref(h)))
        return Batch(self, h)

    def batch_synth_csr(self, n, d_text, seed=42, row0=0):
        h = C.c_void_p()
        check(_lib.lib().twtml_batch_synth_csr(self._h, n, d_text, seed, row0, C.byref(h)))
        return Batch(self, h)


class Batch:
    """One micro-batch resident in HBM (twtml_batch): the RDD[LabeledPoint] of one interval."""

    def __init__(self, ctx, handle):
        self.ctx, self._h = ctx, handle
        n, d, csr = C.c_int64(), C.c_int64(), C.c_int32()
        check(_lib.lib().twtml_batch_info(handle, C.byref(n), C.byref(d), None, C.byref(csr)))
        self.n, self.d, self.is_csr = n.value, d.value, bool(csr.value)
        self._nnz = None
        ctx._children.add(self)

    @property
    def nnz(self):
        """Stored entries; for an asynchronously featurised batch this waits for the featuriser."""
        if self._nnz is None:
            v = C.c_int64()
            check(_lib.lib().twtml_batch_info(self._h, None, None, C.byref(v), None))
            self._nnz = v.value
        return self._nnz

    def count(self):
        return self.n

    def append(self, X, labels):
        """Next rows of a batch from Context.batch_dense_begin (twtml_batch_dense_append)."""
        X = np.asarray(X, np.float64)
        labels = np.ascontiguousarray(labels, np.float64)
        if X.ndim != 2 or X.shape[0] != labels.shape[0] or X.shape[1] != self.d or X.strides[1] != 8 or X.strides[0] % 8:
            raise ValueError("X must be nrows x d float64 with unit column stride and one label per row")
        check(_lib.lib().twtml_batch_dense_append(self._h, X.shape[0], ptr(X), X.strides[0] // 8 if X.shape[0] > 1 else X.shape[1], ptr(labels)))
        self._keep.append((X, labels))
        return self

    def ready(self, wait=False):
        r = C.c_int32()
        check(_lib.lib().twtml_batch_ready(self._h, int(wait), C.byref(r)))
        return bool(r.value)

    def isEmpty(self):
        return self.n == 0

    def export_csr(self):
        rowptr = np.zeros(self.n + 1, np.int64)
        colidx = np.zeros(max(self.nnz, 1), np.int32)
        vals = np.zeros(max(self.nnz, 1), np.float64)
        labels = np.zeros(max(self.n, 1), np.float64)
        check(_lib.lib().twtml_batch_export_csr(self._h, ptr(rowptr), ptr(colidx), ptr(vals), ptr(labels)))
        return rowptr, colidx[:self.nnz], vals[:self.nnz], labels[:self.n]

    def export_dense(self):
        X = np.zeros((self.n, self.d), np.float64)
        labels = np.zeros(max(self.n, 1), np.float64)
        check(_lib.lib().twtml_batch_export_dense(self._h, ptr(X), ptr(labels)))
        return X, labels[:self.n]

    def free(self):
        if self._h:
            _lib.lib().twtml_batch_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


# ------------------------------------------------------------------------------------------------
# org.apache.spark.mllib.linalg / regression value types (data holders only)
# ------------------------------------------------------------------------------------------------
class DenseVector:
    def __init__(self, values):
        self.values = np.asarray(values, np.float64)
        self.size = self.values.shape[0]

    def toArray(self):
        return self.values.copy()


class SparseVector:
    def __init__(self, size, indices, values):
        self.size = int(size)
        self.indices = np.asarray(indices, np.int32)
        self.values = np.asarray(values, np.float64)
        if self.indices.shape != self.values.shape:
            raise ValueError("indices and values differ in length")
        if self.indices.size and (np.any(np.diff(self.indices) <= 0) or self.indices[0] < 0 or self.indices[-1] >= size):
            raise ValueError("SparseVector indices must be strictly increasing and < size")  # Vectors.sparse require() [U]

    def toArray(self):
        a = np.zeros(self.size)
        a[self.indices] = self.values
        return a


class Vectors:
    @staticmethod
    def dense(*values):
        if len(values) == 1 and not np.isscalar(values[0]):
            values = values[0]
        return DenseVector(values)

    @staticmethod
    def sparse(size, indices, values):
        return SparseVector(size, indices, values)

    @staticmethod
    def zeros(size):
        return DenseVector(np.zeros(size))


class LabeledPoint:
    def __init__(self, label, features):
        self.label = float(label)
        self.features = features


def _pack_points(ctx, points, d=None):
    """A list of LabeledPoint (or bare vectors with label 0) -> device batch, CSR if all sparse."""
    pts = [p if isinstance(p, LabeledPoint) else LabeledPoint(0.0, p) for p in points]
    if d is None:
        d = pts[0].features.size if pts else 1
    labels = np.array([p.label for p in pts], np.float64)
    if pts and all(isinstance(p.features, SparseVector) for p in pts):
        rowptr = np.cumsum([0] + [p.features.indices.size for p in pts]).astype(np.int64)
        colidx = np.concatenate([p.features.indices for p in pts]) if pts else np.zeros(0, np.int32)
        vals = np.concatenate([p.features.values for p in pts]) if pts else np.zeros(0)
        return ctx.batch_from_csr(len(pts), d, rowptr, colidx, vals, labels)
    X = np.zeros((len(pts), d), np.float64)
    for i, p in enumerate(pts):
        X[i] = p.features.toArray()
    return ctx.batch_from_dense(X, labels)


# ------------------------------------------------------------------------------------------------
# DStream stand-in
# ------------------------------------------------------------------------------------------------
class DStream:
    """Micro-batch stream: output operations registered with foreachRDD run once per pushed batch, in
    registration order (Spark Streaming's JobScheduler with concurrentJobs = 1).  An "RDD" is whatever
    the producer pushes: a Batch, a list of LabeledPoint, or a list of (key, Vector)."""

    def __init__(self, parent=None, fn=None):
        self._ops, self._children, self._parent, self._fn = [], [], parent, fn

    def foreachRDD(self, fn):
        root = self
        while root._parent is not None:
            root = root._parent
        root._ops.append((self, fn))

    def map(self, fn):
        ch = DStream(self, fn)
        self._children.append(ch)
        return ch

    transform = map

    def cache(self):
        return self

    def _resolve(self, rdd):
        if self._parent is None:
            return rdd
        return self._fn(self._parent._resolve(rdd))

    def push(self, rdd):
        """Deliver one batch interval's RDD to the root stream."""
        for stream, fn in list(self._ops):
            fn(stream._resolve(rdd))


# ------------------------------------------------------------------------------------------------
# the model classes
# ------------------------------------------------------------------------------------------------
class LinearRegressionModel:
    """org.apache.spark.mllib.regression.LinearRegressionModel [U]: weights + intercept (0.0 here)."""

    def __init__(self, owner, weights):
        self._owner = owner
        self.weights = DenseVector(weights)
        self.intercept = 0.0

    def predict(self, x):
        """predict(Vector): Double, or predict(batch / list of vectors): array -- on the GPU."""
        if isinstance(x, (DenseVector, SparseVector)):
            return float(self._owner._predict_batch(_pack_points(self._owner.ctx, [x], self.weights.size))[0])
        if isinstance(x, Batch):
            return self._owner._predict_batch(x)
        return self._owner._predict_batch(_pack_points(self._owner.ctx, list(x), self.weights.size))


class StreamingLinearRegressionWithSGD:
    """Drop-in for org.apache.spark.mllib.regression.StreamingLinearRegressionWithSGD [U] as the
    reference uses it (LinearRegression.scala:28-32,57,86); defaults as in MLlib 1.6.1."""

    def __init__(self, stepSize=0.1, numIterations=50, miniBatchFraction=1.0, ctx=None):
        self.ctx = ctx if ctx is not None else Context(0)
        self.ctx._children.add(self)
        self._m = None
        self._params = dict(step=stepSize, iters=numIterations, frac=miniBatchFraction, tol=0.001, reg=0.0, updater=0)
        self._w0 = None
        self.lastLossHistory = np.zeros(0)
        self.lastStats = None

    def free(self):
        if self._m is not None:
            _lib.lib().twtml_model_free(self._m)
            self._m = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    # ---- setters (return this.type)
    def setStepSize(self, v):
        self._params["step"] = float(v)
        return self._push()

    def setNumIterations(self, v):
        self._params["iters"] = int(v)
        return self._push()

    def setMiniBatchFraction(self, v):
        self._params["frac"] = float(v)
        return self._push()

    def setConvergenceTol(self, v):
        self._params["tol"] = float(v)
        return self._push()

    def setRegParam(self, v):
        self._params["reg"] = float(v)
        return self._push()

    def setUpdater(self, kind):
        self._params["updater"] = {"simple": 0, "l2": 1, "l1": 2}.get(kind, kind)
        return self._push()

    def setInitialWeights(self, v):
        w = v.toArray() if hasattr(v, "toArray") else np.asarray(v, np.float64)
        self._w0 = np.ascontiguousarray(w, np.float64)
        L = _lib.lib()
        if self._m is not None:
            L.twtml_model_free(self._m)
        self._m = C.c_void_p()
        check(L.twtml_model_create(self.ctx._h, self._w0.size, C.byref(self._m)))
        check(L.twtml_set_initial_weights(self._m, ptr(self._w0), self._w0.size))
        return self._push()

    def _push(self):
        if self._m is not None:
            L, p = _lib.lib(), self._params
            check(L.twtml_set_step_size(self._m, p["step"]))
            check(L.twtml_set_num_iterations(self._m, p["iters"]))
            check(L.twtml_set_mini_batch_fraction(self._m, p["frac"]))
            check(L.twtml_set_convergence_tol(self._m, p["tol"]))
            check(L.twtml_set_reg_param(self._m, p["reg"]))
            check(L.twtml_set_updater(self._m, p["updater"]))
        return self

    def _require(self, what):
        if self._m is None:
            # StreamingLinearAlgorithm [U]: IllegalArgumentException
            raise ValueError("Model must be initialized before starting %s." % what)

    # ---- model access
    def latestModel(self):
        self._require("prediction")
        w = np.zeros(self._w0.size)
        check(_lib.lib().twtml_get_weights(self._m, ptr(w), w.size))
        return LinearRegressionModel(self, w)

    def _as_batch(self, rdd):
        if isinstance(rdd, Batch):
            return rdd
        return _pack_points(self.ctx, list(rdd), self._w0.size)

    def _predict_batch(self, batch):
        out = np.zeros(max(batch.n, 1))
        try:
            check(_lib.lib().twtml_predict(self._m, batch._h, ptr(out)))
        except TwtmlError as e:
            raise ValueError(e.message) if e.code == _lib.E_NOT_INITIALIZED else e
        return out[:batch.n]

    # ---- per-batch operations (what the registered output ops call)
    def trainOnBatch(self, rdd):
        """algorithm.run(rdd, model.weights) for one non-empty RDD (StreamingLinearAlgorithm.trainOn [U])."""
        self._require("training")
        b = self._as_batch(rdd)
        hist = np.zeros(max(self._params["iters"], 1))
        it = C.c_int32()
        check(_lib.lib().twtml_train_on_batch(self._m, b._h, ptr(hist), hist.size, C.byref(it)))
        self.lastLossHistory = hist[:it.value].copy()
        return self

    def predictStats(self, rdd, want_arrays=True):
        """The stats output op of LinearRegression.scala:53-66 -> (stats dict, real[], pred[])."""
        self._require("prediction")
        b = self._as_batch(rdd)
        real = np.zeros(max(b.n, 1)) if want_arrays else None
        pred = np.zeros(max(b.n, 1)) if want_arrays else None
        st = _lib.Stats()
        check(_lib.lib().twtml_predict_stats(self._m, b._h, ptr(real), ptr(pred), C.byref(st)))
        self.lastStats = st.as_dict()
        if st.nonfinite:
            raise ArithmeticError("NumberFormatException: Utils.round on a non-finite prediction")  # Utils.scala:4-5
        return self.lastStats, (real[:b.n] if want_arrays else None), (pred[:b.n] if want_arrays else None)

    def predictStatsStatic(self, rdd):
        """The statistics of predictStats, computed without allocating GPU memory (fixed scratch inside
        the library; dense single-GPU batches of at most ~8.4 million rows).  For hosts whose own pool owns
        the GPU, or that cannot take a cudaMalloc stall inside a batch interval."""
        self._require("prediction")
        b = self._as_batch(rdd)
        st = _lib.Stats()
        check(_lib.lib().twtml_predict_stats_static(self._m, b._h, C.byref(st)))
        return st.as_dict()

    def step(self, rdd, want_arrays=True):
        """Both output ops of one interval in one pass: stats with the pre-training weights, then train."""
        self._require("training")
        b = self._as_batch(rdd)
        real = np.zeros(max(b.n, 1)) if want_arrays else None
        pred = np.zeros(max(b.n, 1)) if want_arrays else None
        st = _lib.Stats()
        hist = np.zeros(max(self._params["iters"], 1))
        it = C.c_int32()
        check(_lib.lib().twtml_step(self._m, b._h, ptr(real), ptr(pred), C.byref(st), ptr(hist), hist.size, C.byref(it)))
        self.lastLossHistory = hist[:it.value].copy()
        self.lastStats = st.as_dict()
        return self.lastStats, (real[:b.n] if want_arrays else None), (pred[:b.n] if want_arrays else None)

    # ---- DStream API
    def trainOn(self, stream):
        self._require("training")
        stream.foreachRDD(lambda rdd: None if _is_empty(rdd) else self.trainOnBatch(rdd))

    def predictOn(self, stream):
        self._require("prediction")
        return stream.map(lambda rdd: self._predict_batch(self._as_batch(rdd)))

    def predictOnValues(self, stream):
        self._require("prediction")

        def f(rdd):
            pairs = list(rdd)
            preds = self._predict_batch(_pack_points(self.ctx, [v for _, v in pairs], self._w0.size)) if pairs else []
            return [(k, float(p)) for (k, _), p in zip(pairs, preds)]

        return stream.map(f)


# ------------------------------------------------------------------------------------------------
# the reference's second app (KMeans.scala): StandardScaler + StreamingKMeans, mirrored name for name
# ------------------------------------------------------------------------------------------------
class _Standardized:
    """What StandardScalerModel.transform(rdd) returns in the reference (KMeans.scala:96): the batch's points together with
    the request to scale them by the fitted column deviations.  The scaling itself runs on the GPU inside the model
    update that consumes it; afterwards toArray() gives the scaled points (scaledData.map(_.apply(0)).toArray, :100-101)."""

    def __init__(self, X):
        self.X = np.ascontiguousarray(X, np.float64)
        self.scaled = None
        self.std = None

    def count(self):
        return self.X.shape[0]

    def toArray(self):
        if self.scaled is None:
            raise ValueError("the scaled points exist once a StreamingKMeansModel.update has consumed this batch")
        return self.scaled


class StandardScaler:
    """org.apache.spark.mllib.feature.StandardScaler [U] as KMeans.scala:96 uses it: withMean = false, withStd = true."""

    def __init__(self, withMean=False, withStd=True):
        if withMean or not withStd:
            raise ValueError("only StandardScaler(false, true), the reference's configuration, is offered")

    def fit(self, rdd):
        return self                                      # the fit is part of the GPU call that consumes the batch

    def transform(self, rdd):
        return _Standardized(rdd)


class StreamingKMeansModel:
    """org.apache.spark.mllib.clustering.StreamingKMeansModel [U]: clusterCenters, clusterWeights, update, predict."""

    def __init__(self, owner):
        self._owner = owner

    def _get(self):
        o = self._owner
        c, w = np.zeros((o._k, o._dim)), np.zeros(o._k)
        try:
            check(_lib.lib().twtml_kmeans_get_centers(o._m, ptr(c), ptr(w)))
        except TwtmlError as e:
            raise ValueError(e.message) if e.code == _lib.E_NOT_INITIALIZED else e
        return c, w

    @property
    def clusterCenters(self):
        return self._get()[0]

    @property
    def clusterWeights(self):
        return self._get()[1]

    def update(self, data, decayFactor, timeUnit):
        """update(data, decayFactor, timeUnit) (KMeans.scala:98): data is an n x dim array, or what
        StandardScaler(false, true).fit(rdd).transform(rdd) returned."""
        o = self._owner
        o._require("training")
        L = _lib.lib()
        check(L.twtml_kmeans_set_decay_factor(o._m, float(decayFactor)))
        check(L.twtml_kmeans_set_time_unit(o._m, {"batches": 0, "points": 1}[timeUnit]))
        std = isinstance(data, _Standardized)
        X = data.X if std else np.ascontiguousarray(data, np.float64)
        n = X.shape[0]
        scaled = np.zeros((max(n, 1), o._dim)) if std else None
        sd = np.zeros(o._dim) if std else None
        check(L.twtml_kmeans_update(o._m, ptr(X), n, X.shape[1] if n else o._dim, int(std), ptr(sd), ptr(scaled), None))
        if std:
            data.scaled, data.std = scaled[:n], sd
        return self

    def predict(self, data):
        o = self._owner
        o._require("predictions")
        X = data.toArray() if isinstance(data, _Standardized) else np.ascontiguousarray(data, np.float64)
        out = np.zeros(max(X.shape[0], 1), np.int32)
        check(_lib.lib().twtml_kmeans_predict(o._m, ptr(X), X.shape[0], X.shape[1] if X.shape[0] else o._dim, ptr(out)))
        return out[:X.shape[0]]


class StreamingKMeans:
    """org.apache.spark.mllib.clustering.StreamingKMeans [U] as KMeans.scala:69-73 builds it."""
    BATCHES, POINTS = "batches", "points"

    def __init__(self, k=2, decayFactor=1.0, timeUnit="batches", ctx=None):
        self.ctx = ctx if ctx is not None else Context(0)
        self.ctx._children.add(self)
        self._k, self._dim, self._m = int(k), None, None
        self.decayFactor, self.timeUnit = float(decayFactor), timeUnit

    def free(self):
        if self._m is not None:
            _lib.lib().twtml_kmeans_free(self._m)
            self._m = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def setK(self, k):
        self._k = int(k)
        return self

    def setDecayFactor(self, a):
        self.decayFactor = float(a)
        return self

    def setHalfLife(self, halfLife, timeUnit):
        if timeUnit not in ("batches", "points"):
            raise ValueError("timeUnit must be 'batches' or 'points'")
        self.decayFactor = float(np.exp(np.log(0.5) / halfLife))
        self.timeUnit = timeUnit
        return self

    def _make(self, dim):
        self.free()
        self._dim = int(dim)
        self._m = C.c_void_p()
        check(_lib.lib().twtml_kmeans_create(self.ctx._h, self._k, self._dim, C.byref(self._m)))

    def setInitialCenters(self, centers, weights):
        centers = np.ascontiguousarray(centers, np.float64)
        weights = np.ascontiguousarray(weights, np.float64)
        if centers.shape[0] != self._k or weights.shape != (self._k,):
            raise ValueError("need k centers and k weights")
        self._make(centers.shape[1])
        check(_lib.lib().twtml_kmeans_set_centers(self._m, ptr(centers), ptr(weights)))
        return self

    def setRandomCenters(self, dim, weight, seed=0):
        self._make(dim)
        check(_lib.lib().twtml_kmeans_set_random_centers(self._m, float(weight), int(seed)))
        return self

    def _require(self, what):
        if self._m is None:     # StreamingKMeans.assertInitialized [U]
            raise ValueError("Initial cluster centers must be set before starting %s" % what)

    def latestModel(self):
        self._require("predictions")
        return StreamingKMeansModel(self)

    def trainOn(self, stream):
        self._require("training")
        stream.foreachRDD(lambda rdd: self.latestModel().update(rdd, self.decayFactor, self.timeUnit))

    def predictOn(self, stream):
        self._require("predictions")
        return stream.map(lambda rdd: self.latestModel().predict(rdd))


def stats_payload(count, stats):
    """The JSON body SessionStats.update posts to twtml-web for one batch (SessionStats.scala:22-33 ->
    web/WebClient.scala:21-38, ApiTypes.scala:5-17; json4s ShortTypeHints put the class name under "jsonClass"):
    all five numbers are Longs -- the already Utils.round-ed statistics truncated with .toLong."""
    import json
    return json.dumps({"jsonClass": "Stats", "count": int(count), "batch": int(stats["batch"]), "mse": int(stats["mse"]),
                       "realStddev": int(stats["real_stdev"]), "predStddev": int(stats["pred_stdev"])}, separators=(",", ":"))


def lightning_series(stats, real, pred):
    """The four series SessionStats.update appends to the Lightning line-streaming plot for one batch
    (SessionStats.scala:26-33): Array(real, pred, Array.fill(batch)(realStdev), Array.fill(batch)(predStdev)) --
    the labels, the Utils.round-ed predictions, and the two already rounded standard deviations repeated once
    per row (batch.toInt of them).  `stats`, `real`, `pred` are what predictStats / step return."""
    n = int(stats["batch"])
    return [np.asarray(real, np.float64), np.asarray(pred, np.float64),
            np.full(n, float(stats["real_stdev"])), np.full(n, float(stats["pred_stdev"]))]


def _is_empty(rdd):
    return rdd.isEmpty() if isinstance(rdd, Batch) else len(rdd) == 0


# ------------------------------------------------------------------------------------------------
# MllibHelper (MllibHelper.scala) over batches of tweets
# ------------------------------------------------------------------------------------------------
class TweetBatch:
    """Struct-of-arrays form of the twitter4j.Status fields the featuriser reads
    (MllibHelper.scala:43-45,59-61,81,85,91).  Text is the retweeted status text as UTF-16 code
    units; lower-casing (toLowerCase, :45) is the caller's job unless lowercase=True is passed on."""

    def __init__(self, texts, retweet_count, followers, favourites, friends, created_ms, is_retweet):
        units = [np.frombuffer(t.encode("utf-16-le", "surrogatepass"), np.uint16) if isinstance(t, str)
                 else np.asarray(t, np.uint16) for t in texts]
        self.text_off = np.cumsum([0] + [u.size for u in units]).astype(np.int64)
        self.text_utf16 = np.concatenate(units).astype(np.uint16) if units and self.text_off[-1] else np.zeros(0, np.uint16)
        self.retweet_count = np.ascontiguousarray(retweet_count, np.int32)
        self.followers = np.ascontiguousarray(followers, np.int32)
        self.favourites = np.ascontiguousarray(favourites, np.int32)
        self.friends = np.ascontiguousarray(friends, np.int32)
        self.created_ms = np.ascontiguousarray(created_ms, np.int64)
        self.is_retweet = np.ascontiguousarray(is_retweet, np.uint8)

    @classmethod
    def from_arrays(cls, text_utf16, text_off, retweet_count, followers, favourites, friends, created_ms, is_retweet):
        """Already packed columns (text as one uint16 array + n+1 offsets)."""
        self = cls.__new__(cls)
        self.text_utf16 = np.ascontiguousarray(text_utf16, np.uint16)
        self.text_off = np.ascontiguousarray(text_off, np.int64)
        self.retweet_count = np.ascontiguousarray(retweet_count, np.int32)
        self.followers = np.ascontiguousarray(followers, np.int32)
        self.favourites = np.ascontiguousarray(favourites, np.int32)
        self.friends = np.ascontiguousarray(friends, np.int32)
        self.created_ms = np.ascontiguousarray(created_ms, np.int64)
        self.is_retweet = np.ascontiguousarray(is_retweet, np.uint8)
        return self

    def __len__(self):
        return int(self.retweet_count.shape[0])


class MllibHelper:
    """MllibHelper.scala: the vars of :13-20 and reset() with the reference's shadowing quirk --
    reset() (:22-32) re-declares hashText / numFeatures / numberFeatureIndices as locals, so a
    configured numTextFeatures never reaches the featuriser: it always hashes into 1000 buckets
    and the feature width stays 1004."""
    numNumberFeatures = 4
    numRetweetBegin = 100
    numRetweetEnd = 1000
    numTextFeatures = 1000
    _hashTextFeatures = 1000      # `hashText = new HashingTF(numTextFeatures)` at object init (:18)
    numFeatures = 1004            # :19, never reassigned (shadowed at :28)

    @classmethod
    def reset(cls, numRetweetBegin=100, numRetweetEnd=1000, numTextFeatures=1000):
        cls.numRetweetBegin, cls.numRetweetEnd, cls.numTextFeatures = numRetweetBegin, numRetweetEnd, numTextFeatures
        # :27-29 assign locals only: _hashTextFeatures and numFeatures keep their initial values

    @classmethod
    def featurize_batch(cls, ctx, tweets, now_ms, lowercase=False):
        """stream.filter(filtrate).map(featurize) for one interval's tweets -> CSR Batch on the GPU."""
        return ctx.batch_from_tweets(tweets, now_ms, cls.numRetweetBegin, cls.numRetweetEnd, cls._hashTextFeatures, lowercase)
