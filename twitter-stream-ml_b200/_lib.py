"""ctypes binding of libtwtml.so (include/twtml.h) and the nvcc recipe that builds it in-tree.

This is the only way into the product's arithmetic: if the shared library is missing or fails to
load, importing a handle raises -- there is no Python or CPU fallback for any entry point.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO_PATH = os.path.join(HERE, "libtwtml.so")
HEADER = os.path.join(os.path.dirname(HERE), "include", "twtml.h")
SOURCES = ["twtml.cu", "twtml_sparse.cu", "twtml_noalloc.cu", "twtml_group.cu", "twtml_kmeans.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared"]

OK, E_INVALID_ARG, E_NOT_INITIALIZED, E_CUDA, E_NCCL, E_OOM, E_SHAPE_MISMATCH, E_NONFINITE = 0, -1, -2, -3, -4, -5, -6, -7
NCCL_ID_BYTES = 128
FLAG_LOWERCASE = 1


class TwtmlError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libtwtml error %d: %s" % (code, msg))
        self.code = code
        self.message = msg


class Stats(C.Structure):
    _fields_ = [("batch", C.c_int64), ("mse", C.c_double), ("real_stdev", C.c_double),
                ("pred_stdev", C.c_double), ("mse_raw", C.c_double), ("real_stdev_raw", C.c_double),
                ("pred_stdev_raw", C.c_double), ("real_mean", C.c_double), ("pred_mean", C.c_double),
                ("nonfinite", C.c_int64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class Profile(C.Structure):
    _fields_ = [("iter_launches", C.c_int32), ("other_launches", C.c_int32), ("nccl_calls", C.c_int32),
                ("iter_kernel_ms", C.c_float), ("loop_ms", C.c_float), ("iter_kernel_bytes", C.c_int64)]


def _stale():
    if not os.path.exists(SO_PATH):
        return True
    t = os.path.getmtime(SO_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [HEADER]
    return any(os.path.getmtime(p) > t for p in deps)


def build(force=False, verbose=False):
    """Compile libtwtml.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    if not force and not _stale():
        return SO_PATH
    cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", SO_PATH] + [os.path.join(CSRC, s) for s in SOURCES] + ["-ldl"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return SO_PATH


_lib = None
_vp, _i32, _i64, _u64, _f64, _u32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_double, C.c_uint32

# name -> argtypes; every function returns int32 status.  Mirrors include/twtml.h one to one.
SIGNATURES = {
    "twtml_abi_version": [],
    "twtml_last_error": [C.c_char_p, _i64],
    "twtml_create": [_i32, C.POINTER(_vp)],
    "twtml_nccl_unique_id": [_vp, _i64],
    "twtml_create_ranked": [_i32, _i32, _i32, _vp, _i64, C.POINTER(_vp)],
    "twtml_create_all": [_vp, _i32, C.POINTER(_vp)],
    "twtml_device_count": [_vp, C.POINTER(_i32)],
    "twtml_destroy": [_vp],
    "twtml_rank": [_vp, C.POINTER(_i32), C.POINTER(_i32)],
    "twtml_profile_enable": [_vp, _i32],
    "twtml_profile_get": [_vp, C.POINTER(Profile)],
    "twtml_model_create": [_vp, _i64, C.POINTER(_vp)],
    "twtml_model_free": [_vp],
    "twtml_set_step_size": [_vp, _f64],
    "twtml_set_num_iterations": [_vp, _i32],
    "twtml_set_mini_batch_fraction": [_vp, _f64],
    "twtml_set_convergence_tol": [_vp, _f64],
    "twtml_set_reg_param": [_vp, _f64],
    "twtml_set_updater": [_vp, _i32],
    "twtml_set_initial_weights": [_vp, _vp, _i64],
    "twtml_get_weights": [_vp, _vp, _i64],
    "twtml_num_features": [_vp, C.POINTER(_i64)],
    "twtml_batch_from_tweets": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i32, _i32, _i32, _u32, C.POINTER(_vp)],
    "twtml_batch_from_csr": [_vp, _i64, _i64, _vp, _vp, _vp, _vp, C.POINTER(_vp)],
    "twtml_batch_from_dense": [_vp, _i64, _i64, _vp, _i64, _vp, C.POINTER(_vp)],
    "twtml_batch_from_dense_async": [_vp, _i64, _i64, _vp, _i64, _vp, C.POINTER(_vp)],
    "twtml_batch_dense_begin": [_vp, _i64, _i64, C.POINTER(_vp)],
    "twtml_batch_dense_append": [_vp, _i64, _vp, _i64, _vp],
    "twtml_batch_from_csr_async": [_vp, _i64, _i64, _vp, _vp, _vp, _vp, C.POINTER(_vp)],
    "twtml_batch_from_tweets_async": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i32, _i32, _i32, _u32, C.POINTER(_vp)],
    "twtml_batch_ready": [_vp, _i32, C.POINTER(_i32)],
    "twtml_batch_synth_dense": [_vp, _i64, _i64, _u64, _i64, C.POINTER(_vp)],
    "twtml_batch_synth_csr": [_vp, _i64, _i32, _u64, _i64, C.POINTER(_vp)],
    "twtml_batch_rows": [_vp, C.POINTER(_i64)],
    "twtml_batch_info": [_vp, C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i32)],
    "twtml_batch_export_csr": [_vp, _vp, _vp, _vp, _vp],
    "twtml_batch_export_dense": [_vp, _vp, _vp],
    "twtml_batch_free": [_vp],
    "twtml_predict": [_vp, _vp, _vp],
    "twtml_predict_stats": [_vp, _vp, _vp, _vp, C.POINTER(Stats)],
    "twtml_predict_stats_static": [_vp, _vp, C.POINTER(Stats)],
    "twtml_train_on_batch": [_vp, _vp, _vp, _i32, C.POINTER(_i32)],
    "twtml_step": [_vp, _vp, _vp, _vp, C.POINTER(Stats), _vp, _i32, C.POINTER(_i32)],
    "twtml_kmeans_create": [_vp, _i32, _i32, C.POINTER(_vp)],
    "twtml_kmeans_free": [_vp],
    "twtml_kmeans_set_decay_factor": [_vp, _f64],
    "twtml_kmeans_set_time_unit": [_vp, _i32],
    "twtml_kmeans_set_half_life": [_vp, _f64, _i32],
    "twtml_kmeans_set_centers": [_vp, _vp, _vp],
    "twtml_kmeans_set_random_centers": [_vp, _f64, _u64],
    "twtml_kmeans_get_centers": [_vp, _vp, _vp],
    "twtml_kmeans_update": [_vp, _vp, _i64, _i64, _i32, _vp, _vp, _vp],
    "twtml_kmeans_predict": [_vp, _vp, _i64, _i64, _vp],
}


def lib():
    """Load libtwtml.so (building it first when nvcc is available and sources are newer)."""
    global _lib
    if _lib is None:
        if _stale():
            try:
                build()
            except FileNotFoundError:
                pass  # no nvcc on this host: use the shipped .so if there is one
        if not os.path.exists(SO_PATH):
            raise ImportError("libtwtml.so is not built (%s); run __graft_entry__.build() -- there is no fallback path" % SO_PATH)
        L = C.CDLL(SO_PATH)
        for name, args in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = _i32
            fn.argtypes = args
        _lib = L
    return _lib


def last_error():
    buf = C.create_string_buffer(512)
    lib().twtml_last_error(buf, 512)
    return buf.value.decode("utf-8", "replace")


def check(rc):
    if rc != 0:
        raise TwtmlError(rc, last_error())


def ptr(a):
    return None if a is None else a.ctypes.data_as(_vp)


def arr(a, dt):
    return np.ascontiguousarray(a, dtype=dt)
