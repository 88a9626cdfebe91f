"""ctypes front-end to the C oracle (oracle/dq4ml_oracle.c).

TEST INFRASTRUCTURE ONLY -- imported by tests/, __graft_entry__.smoke() and bench.py's CPU
baseline legs, never by the product package.  PARITY UNPINNED by the reference's own tests
(it has none); see dq4ml_oracle.h.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libdq4ml_oracle.so")

ORC_I32, ORC_F64 = 0, 1


class _Column(C.Structure):
    _fields_ = [("data", C.c_void_p), ("dtype", C.c_int32)]


class SynthParams(C.Structure):
    _fields_ = [
        ("seed_data", C.c_uint64),
        ("seed_anom", C.c_uint64),
        ("p_low", C.c_double),
        ("p_high", C.c_double),
        ("d", C.c_int32),
    ]


def build(force: bool = False) -> str:
    """Compile the C oracle in place (gcc, seconds)."""
    src = os.path.join(_HERE, "dq4ml_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"], env={**os.environ, "CC": "/usr/bin/gcc"})
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        dp, ip, u8p = C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_uint8)
        L.orc_check_minimum_price.restype = C.c_double
        L.orc_check_minimum_price.argtypes = [C.c_double]
        L.orc_check_price_range.restype = C.c_double
        L.orc_check_price_range.argtypes = [C.c_double, C.c_int32]
        L.orc_spark_gt_zero.argtypes = [C.c_double]
        L.orc_dq_keep.argtypes = [C.c_int32, C.c_double]
        L.orc_udf_minimum_price.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.orc_udf_price_correlation.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.orc_udf_price_correlation.restype = None
        L.orc_dq_filter.restype = C.c_int64
        L.orc_dq_filter.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]
        L.orc_csv_parse.restype = C.c_int64
        L.orc_csv_parse.argtypes = [C.c_char_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64]
        colp = C.POINTER(_Column)
        L.orc_wls_moments.restype = None
        L.orc_wls_moments.argtypes = [colp, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_summarizer.restype = None
        L.orc_summarizer.argtypes = [colp, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_lsq_loss_grad.restype = None
        L.orc_lsq_loss_grad.argtypes = [colp, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_int, C.c_void_p]
        L.orc_predict_metrics.restype = None
        L.orc_predict_metrics.argtypes = [colp, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_double, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_synth_generate.restype = None
        L.orc_synth_generate.argtypes = [C.POINTER(SynthParams), C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_int64, C.c_int]
        L.orc_rows_digest.restype = None
        L.orc_rows_digest.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_uint64, C.c_void_p]
        L.orc_max_threads.restype = C.c_int
        L.orc_lab_rowwise_moments.restype = None
        L.orc_lab_rowwise_moments.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p]
        _lib = L
    return _lib


def max_threads() -> int:
    """What OpenMP would use by default.  Under torchrun this is 1 (it exports OMP_NUM_THREADS=1): callers that want the
    machine use host_threads()."""
    return int(lib().orc_max_threads())


def host_threads() -> int:
    """Hardware threads this process may run on (its CPU affinity mask), independent of OMP_NUM_THREADS -- the thread
    count of Spark's local[*] master (DataQuality4MachineLearningApp.java:40)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except (AttributeError, OSError):
        return max(1, os.cpu_count() or 1)


def _ptr(a):
    return None if a is None else a.ctypes.data


def _cols(features):
    """features: list of 1-D numpy arrays (int32 or float64), all length n."""
    arr = (_Column * max(len(features), 1))()
    keep = []
    for j, f in enumerate(features):
        if f.dtype == np.int32:
            dt = ORC_I32
        elif f.dtype == np.float64:
            dt = ORC_F64
        else:
            raise TypeError(f"feature column {j}: dtype {f.dtype} not in (int32, float64)")
        f = np.ascontiguousarray(f)
        keep.append(f)
        arr[j].data = f.ctypes.data
        arr[j].dtype = dt
    return arr, keep


def _dq(dq):
    if dq is None:
        return None, None, ()
    g = np.ascontiguousarray(dq[0], dtype=np.int32)
    p = np.ascontiguousarray(dq[1], dtype=np.float64)
    return g.ctypes.data, p.ctypes.data, (g, p)


# ---------------------------------------------------------------- DQ

def check_minimum_price(price: float) -> float:
    return lib().orc_check_minimum_price(float(price))


def check_price_range(price: float, guest: int) -> float:
    return lib().orc_check_price_range(float(price), int(guest))


def udf_minimum_price(price, valid=None):
    price = np.ascontiguousarray(price, dtype=np.float64)
    out = np.empty_like(price)
    v = None if valid is None else np.ascontiguousarray(valid, dtype=np.uint8)
    rc = lib().orc_udf_minimum_price(price.ctypes.data, _ptr(v), price.size, out.ctypes.data)
    if rc != 0:
        raise ValueError("null price reached minimumPriceRule (NullPointerException in the reference, MinUdf:12)")
    return out


def udf_price_correlation(price, guest, price_valid=None, guest_valid=None):
    price = np.ascontiguousarray(price, dtype=np.float64)
    guest = np.ascontiguousarray(guest, dtype=np.int32)
    pv = None if price_valid is None else np.ascontiguousarray(price_valid, dtype=np.uint8)
    gv = None if guest_valid is None else np.ascontiguousarray(guest_valid, dtype=np.uint8)
    out = np.empty_like(price)
    lib().orc_udf_price_correlation(price.ctypes.data, _ptr(pv), guest.ctypes.data, _ptr(gv), price.size, out.ctypes.data)
    return out


def dq_filter(guest, price):
    guest = np.ascontiguousarray(guest, dtype=np.int32)
    price = np.ascontiguousarray(price, dtype=np.float64)
    go = np.empty_like(guest)
    po = np.empty_like(price)
    n1 = C.c_int64(0)
    m = lib().orc_dq_filter(guest.ctypes.data, price.ctypes.data, guest.size, go.ctypes.data, po.ctypes.data, C.byref(n1))
    return go[:m].copy(), po[:m].copy(), int(n1.value)


def csv_parse(data: bytes):
    cap = len(data) // 2 + 2
    g = np.empty(cap, dtype=np.int32)
    p = np.empty(cap, dtype=np.float64)
    n = lib().orc_csv_parse(data, len(data), g.ctypes.data, p.ctypes.data, cap)
    if n < 0:
        raise ValueError("malformed CSV")
    return g[:n].copy(), p[:n].copy()


# ---------------------------------------------------------------- aggregation passes

def wls_moments(features, label, dq=None, threads=1):
    d = len(features)
    label = np.ascontiguousarray(label, dtype=np.float64)
    cols, keep = _cols(features)
    out = np.zeros(5 + 2 * d + d * (d + 1) // 2, dtype=np.float64)
    g, p, k2 = _dq(dq)
    lib().orc_wls_moments(cols, d, label.ctypes.data, label.size, g, p, threads, out.ctypes.data)
    return out


def lab_rowwise_moments(guest, price, threads=1):
    """The lab's d = 1 pipeline one row at a time through the boxed UDF call chain (the row-at-a-time CPU-baseline
    variant); same layout and values as wls_moments([guest], price, dq=(guest, price))."""
    guest = np.ascontiguousarray(guest, dtype=np.int32)
    price = np.ascontiguousarray(price, dtype=np.float64)
    out = np.zeros(8, dtype=np.float64)
    lib().orc_lab_rowwise_moments(guest.ctypes.data, price.ctypes.data, guest.size, threads, out.ctypes.data)
    return out


def summarizer(features, label, dq=None, threads=1):
    d = len(features)
    label = np.ascontiguousarray(label, dtype=np.float64)
    cols, keep = _cols(features)
    out = np.zeros(3 + 7 * (d + 1), dtype=np.float64)
    g, p, k2 = _dq(dq)
    lib().orc_summarizer(cols, d, label.ctypes.data, label.size, g, p, threads, out.ctypes.data)
    stats = out[3:].reshape(d + 1, 7)
    return {
        "count": int(out[0]), "wsum": out[1], "wwsum": out[2],
        "mean": stats[:, 0].copy(), "variance": stats[:, 1].copy(), "min": stats[:, 2].copy(),
        "max": stats[:, 3].copy(), "nnz": stats[:, 4].copy(), "normL1": stats[:, 5].copy(),
        "normL2": stats[:, 6].copy(),
    }


def lsq_loss_grad(features, label, coef, fstd, fmean, ystd, ymean, fit_intercept=True, dq=None, threads=1):
    d = len(features)
    label = np.ascontiguousarray(label, dtype=np.float64)
    cols, keep = _cols(features)
    coef = np.ascontiguousarray(coef, dtype=np.float64)
    fstd = np.ascontiguousarray(fstd, dtype=np.float64)
    fmean = np.ascontiguousarray(fmean, dtype=np.float64)
    out = np.zeros(2 + d, dtype=np.float64)
    g, p, k2 = _dq(dq)
    lib().orc_lsq_loss_grad(cols, d, label.ctypes.data, label.size, g, p, coef.ctypes.data, fstd.ctypes.data,
                            fmean.ctypes.data, float(ystd), float(ymean), int(bool(fit_intercept)), threads,
                            out.ctypes.data)
    return out[0], out[1], out[2:].copy()


def predict_metrics(features, label, coef, intercept, dq=None, threads=1, want_prediction=True):
    d = len(features)
    n = features[0].size if d else (0 if label is None else np.asarray(label).size)
    cols, keep = _cols(features)
    lab = None if label is None else np.ascontiguousarray(label, dtype=np.float64)
    coef = np.ascontiguousarray(coef, dtype=np.float64)
    pred = np.empty(n, dtype=np.float64) if want_prediction else None
    out = np.zeros(5, dtype=np.float64)
    g, p, k2 = _dq(dq)
    lib().orc_predict_metrics(cols, d, _ptr(lab), n, g, p, coef.ctypes.data, float(intercept), threads, _ptr(pred),
                              out.ctypes.data)
    return pred, out


# ---------------------------------------------------------------- synthetic data

def synth_generate(n, d=1, row0=0, seed_data=20200601, seed_anom=7, p_low=0.01, p_high=0.01, threads=0):
    """Returns guest:int32[n], price:f64[n], extra:f64[d-1, n]."""
    if threads <= 0:
        threads = host_threads()
    sp = SynthParams(seed_data, seed_anom, p_low, p_high, d)
    guest = np.empty(n, dtype=np.int32)
    price = np.empty(n, dtype=np.float64)
    extra = np.empty((max(d - 1, 0), n), dtype=np.float64)
    lib().orc_synth_generate(C.byref(sp), row0, n, guest.ctypes.data, price.ctypes.data,
                             extra.ctypes.data if d > 1 else None, n, threads)
    return guest, price, extra


def rows_digest(guest, price, seed=0):
    guest = np.ascontiguousarray(guest, dtype=np.int32)
    price = np.ascontiguousarray(price, dtype=np.float64)
    out = np.zeros(2, dtype=np.uint64)
    lib().orc_rows_digest(guest.ctypes.data, price.ctypes.data, guest.size, seed, out.ctypes.data)
    return int(out[0]), int(out[1])
