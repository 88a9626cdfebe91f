"""ctypes binding of libdq4ml.so (include/dq4ml.h).  This is the same C ABI the Java drop-in binds through Panama
(INTEGRATION.md); Python is only the harness language here because the image has no JVM.

The library is built in-tree by csrc/Makefile (see __graft_entry__.build).  There is no fallback: if the shared
object is missing, importing this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdq4ml.so")

OK = 0
ERR_INVALID, ERR_CUDA, ERR_NULL_VALUE, ERR_NO_DEVICE, ERR_SINGULAR, ERR_COMM, ERR_UNSUPPORTED, ERR_EMPTY = (
    -1, -2, -3, -4, -5, -6, -7, -8)
I32, F64 = 0, 1
SOLVER_AUTO, SOLVER_NORMAL, SOLVER_LBFGS = 0, 1, 2
SOLVERS = {"auto": SOLVER_AUTO, "normal": SOLVER_NORMAL, "l-bfgs": SOLVER_LBFGS}
LOSSES = {"squaredError": 0, "huber": 1}
ABI_VERSION = 2


class Dq4mlError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"dq4ml status {status}: {message}")
        self.status = status


class NullValueError(Dq4mlError):
    """A null reached minimumPriceRule -- the reference throws NullPointerException (MinimumPriceDataQualityUdf.java:12)."""


class Column(C.Structure):
    _fields_ = [("data", C.c_void_p), ("dtype", C.c_int32)]


class Columns(C.Structure):
    _fields_ = [("cols", C.POINTER(Column)), ("n_cols", C.c_int32)]


class DqParams(C.Structure):
    _fields_ = [("min_price", C.c_double), ("min_guest", C.c_int32), ("max_price", C.c_double)]


class DqFused(C.Structure):
    _fields_ = [("guest", C.c_void_p), ("price", C.c_void_p), ("params", DqParams)]


class LrParams(C.Structure):
    _fields_ = [("max_iter", C.c_int32), ("reg_param", C.c_double), ("elastic_net_param", C.c_double),
                ("tol", C.c_double), ("fit_intercept", C.c_int32), ("standardization", C.c_int32),
                ("solver", C.c_int32), ("loss", C.c_int32), ("epsilon", C.c_double)]


class SynthParams(C.Structure):
    _fields_ = [("seed_data", C.c_uint64), ("seed_anom", C.c_uint64), ("p_low", C.c_double), ("p_high", C.c_double),
                ("d", C.c_int32)]


ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_double), C.c_int64)

# name -> (restype, argtypes); every symbol include/dq4ml.h declares
SIGNATURES = {
    "dq4ml_last_error": (C.c_char_p, []),
    "dq4ml_abi_version": (C.c_int, []),
    "dq4ml_ctx_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "dq4ml_ctx_destroy": (None, [C.c_void_p]),
    "dq4ml_ctx_synchronize": (C.c_int, [C.c_void_p]),
    "dq4ml_ctx_stream": (C.c_void_p, [C.c_void_p]),
    "dq4ml_ctx_launch_count": (C.c_int64, [C.c_void_p]),
    "dq4ml_ctx_enable_pass_timing": (C.c_int, [C.c_void_p, C.c_int]),
    "dq4ml_ctx_pass_timing": (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "dq4ml_ctx_pass_times": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32]),
    "dq4ml_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "dq4ml_ctx_init_nccl": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "dq4ml_ctx_set_allreduce": (C.c_int, [C.c_void_p, ALLREDUCE_FN, C.c_void_p, C.c_int, C.c_int]),
    "dq4ml_ctx_allreduce": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "dq4ml_device_alloc": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]),
    "dq4ml_device_free": (C.c_int, [C.c_void_p, C.c_void_p]),
    "dq4ml_host_alloc_pinned": (C.c_int, [C.c_int64, C.POINTER(C.c_void_p)]),
    "dq4ml_host_free_pinned": (C.c_int, [C.c_void_p]),
    "dq4ml_memcpy": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "dq4ml_dq_params_default": (DqParams, []),
    "dq4ml_check_minimum_price": (C.c_double, [C.c_double]),
    "dq4ml_check_price_range": (C.c_double, [C.c_double, C.c_int32]),
    "dq4ml_udf_minimumPriceRule": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(DqParams)]),
    "dq4ml_udf_priceCorrelationRule": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                                 C.c_void_p, C.POINTER(DqParams)]),
    "dq4ml_dq_filter": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(DqParams), C.c_void_p,
                                  C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "dq4ml_dq_filter_global_offset": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "dq4ml_dq_filter_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(DqParams), C.c_void_p,
                                          C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "dq4ml_csv_parse": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "dq4ml_lr_params_default": (LrParams, []),
    "dq4ml_wls_moments_len": (C.c_int64, [C.c_int32]),
    "dq4ml_wls_moments": (C.c_int, [C.c_void_p, C.POINTER(Columns), C.c_void_p, C.c_int64, C.POINTER(DqFused), C.c_void_p]),
    "dq4ml_wls_solve": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(LrParams), C.POINTER(C.c_void_p)]),
    "dq4ml_summary_len": (C.c_int64, [C.c_int32]),
    "dq4ml_summarize": (C.c_int, [C.c_void_p, C.POINTER(Columns), C.c_void_p, C.c_int64, C.POINTER(DqFused), C.c_void_p]),
    "dq4ml_lsq_loss_grad": (C.c_int, [C.c_void_p, C.POINTER(Columns), C.c_void_p, C.c_int64, C.POINTER(DqFused), C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_void_p]),
    "dq4ml_linreg_fit": (C.c_int, [C.c_void_p, C.POINTER(Columns), C.c_void_p, C.c_int64, C.POINTER(LrParams),
                                   C.POINTER(DqFused), C.POINTER(C.c_void_p)]),
    "dq4ml_ctx_set_gram_tier": (C.c_int, [C.c_void_p, C.c_int]),
    "dq4ml_linreg_fit_weighted": (C.c_int, [C.c_void_p, C.POINTER(Columns), C.c_void_p, C.c_void_p, C.c_int64,
                                            C.POINTER(LrParams), C.POINTER(DqFused), C.POINTER(C.c_void_p)]),
    "dq4ml_weighted_residual_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(DqFused),
                                                C.c_void_p]),
    "dq4ml_linreg_model_num_features": (C.c_int32, [C.c_void_p]),
    "dq4ml_linreg_model_coefficients": (C.c_int, [C.c_void_p, C.c_void_p]),
    "dq4ml_linreg_model_intercept": (C.c_double, [C.c_void_p]),
    "dq4ml_linreg_model_total_iterations": (C.c_int32, [C.c_void_p]),
    "dq4ml_linreg_model_objective_history": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32]),
    "dq4ml_linreg_model_data_passes": (C.c_int32, [C.c_void_p]),
    "dq4ml_linreg_model_scale": (C.c_double, [C.c_void_p]),
    "dq4ml_linreg_model_num_rows": (C.c_int64, [C.c_void_p]),
    "dq4ml_linreg_model_solver_used": (C.c_char_p, [C.c_void_p]),
    "dq4ml_linreg_predict1": (C.c_double, [C.c_void_p, C.c_void_p]),
    "dq4ml_linreg_model_inference": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_int32]),
    "dq4ml_linreg_model_destroy": (None, [C.c_void_p]),
    "dq4ml_linreg_model_create": (C.c_int, [C.c_void_p, C.c_int32, C.c_double, C.POINTER(C.c_void_p)]),
    "dq4ml_linreg_transform": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Columns), C.c_int64, C.c_void_p]),
    "dq4ml_linreg_evaluate": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Columns), C.c_void_p, C.c_int64,
                                        C.POINTER(DqFused), C.c_int, C.c_void_p, C.c_void_p]),
    "dq4ml_synth_generate": (C.c_int, [C.c_void_p, C.POINTER(SynthParams), C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_int64]),
    "dq4ml_rows_digest": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_uint64, C.c_void_p]),
}

_lib = None


def lib():
    """Loads libdq4ml.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU/Python fallback for the CUDA path)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        if L.dq4ml_abi_version() != ABI_VERSION:
            raise ImportError("libdq4ml.so ABI version mismatch")
        _lib = L
    return _lib


def check(status):
    if status < 0:
        msg = lib().dq4ml_last_error().decode(errors="replace")
        if status == ERR_NULL_VALUE:
            raise NullValueError(status, msg)
        raise Dq4mlError(status, msg)
    return status


# ---------------------------------------------------------------------------------------------- buffers

def is_torch(x):
    return type(x).__module__.startswith("torch")


def buf_ptr(x):
    """Address of a numpy array (host) or torch tensor (host or cuda); None -> NULL."""
    if x is None:
        return None
    if is_torch(x):
        if not x.is_contiguous():
            raise ValueError("tensor must be contiguous")
        return x.data_ptr()
    if isinstance(x, np.ndarray):
        if not x.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return x.ctypes.data
    if isinstance(x, int):
        return x
    raise TypeError(f"unsupported buffer type {type(x)}")


def buf_dtype(x):
    s = str(x.dtype)
    if s in ("int32", "torch.int32"):
        return I32
    if s in ("float64", "torch.float64"):
        return F64
    raise TypeError(f"column dtype {s} not in (int32, float64)")


def buf_len(x):
    return int(x.numel()) if is_torch(x) else int(x.size)


def make_columns(cols):
    """cols: sequence of numpy arrays / torch tensors -> (Columns struct, keep-alive tuple)."""
    arr = (Column * max(len(cols), 1))()
    for j, c in enumerate(cols):
        arr[j].data = buf_ptr(c)
        arr[j].dtype = buf_dtype(c)
    tab = Columns(arr, len(cols))
    return tab, (arr, cols)


def make_dq(guest, price, params=None):
    if guest is None:
        return None, ()
    if buf_dtype(guest) != I32 or buf_dtype(price) != F64:
        raise TypeError("fused DQ needs guest:int32 and price:float64")
    p = params if params is not None else lib().dq4ml_dq_params_default()
    return DqFused(buf_ptr(guest), buf_ptr(price), p), (guest, price)
