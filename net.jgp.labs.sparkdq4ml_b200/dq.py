"""Host-side mirror of the lab's DQ plug-in surface, calling the C ABI.

Reference classes mirrored (same names, same argument meaning, same error behaviour):
  dq/service/MinimumPriceDataQualityService.java:3-15      -> MinimumPriceDataQualityService.checkMinimumPrice
  dq/service/PriceCorrelationDataQualityService.java:3-12  -> PriceCorrelationDataQualityService.checkPriceRange
  dq/udf/MinimumPriceDataQualityUdf.java:7-14              -> MinimumPriceDataQualityUdf.call (UDF1<Double,Double>)
  dq/udf/PriceCorrelationDataQualityUdf.java:7-18          -> PriceCorrelationDataQualityUdf.call (UDF2<Double,Integer,Double>)
  DataQuality4MachineLearningApp.java:46-49                -> UDFRegistration.register("minimumPriceRule" | "priceCorrelationRule", ...)
  DataQuality4MachineLearningApp.java:3,69,87              -> callUDF(name, cols...)
  DataQuality4MachineLearningApp.java:76-78,88-90          -> dq_filter (rule -> WHERE col > 0 -> projection)

`call` keeps Spark's row-at-a-time contract (scalars, None = SQL NULL); `call_columns` is the columnar fast path that
runs the sm_100a kernels on whole columns (numpy host arrays or torch cuda tensors).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi as capi

DoubleType = "double"  # stands in for org.apache.spark.sql.types.DataTypes.DoubleType (App:47,49)


class MinimumPriceDataQualityService:
    MIN_PRICE = 20

    @staticmethod
    def checkMinimumPrice(price: float) -> float:
        return capi.lib().dq4ml_check_minimum_price(float(price))


class PriceCorrelationDataQualityService:
    @staticmethod
    def checkPriceRange(price: float, guest: int) -> float:
        return capi.lib().dq4ml_check_price_range(float(price), int(guest))


def _alloc_like(col, dtype_np, dtype_torch_name):
    if capi.is_torch(col):
        import torch
        return torch.empty(col.shape, dtype=getattr(torch, dtype_torch_name), device=col.device)
    return np.empty(col.shape, dtype=dtype_np)


class MinimumPriceDataQualityUdf:
    """UDF1<Double, Double>; serialVersionUID kept for the Java drop-in (MinimumPriceDataQualityUdf.java:9)."""

    serialVersionUID = -201966159201746851
    name = "minimumPriceRule"

    def call(self, t1):
        if t1 is None:  # Java auto-unboxing of a null Double (MinUdf:12)
            raise capi.NullValueError(capi.ERR_NULL_VALUE, "NullPointerException: null price in minimumPriceRule")
        return MinimumPriceDataQualityService.checkMinimumPrice(t1)

    def call_columns(self, ctx, price, price_valid=None, params=None):
        out = _alloc_like(price, np.float64, "float64")
        p = C.byref(params) if params is not None else None
        capi.check(capi.lib().dq4ml_udf_minimumPriceRule(ctx.handle, capi.buf_ptr(price), capi.buf_ptr(price_valid),
                                                          capi.buf_len(price), capi.buf_ptr(out), p))
        return out


class PriceCorrelationDataQualityUdf:
    """UDF2<Double, Integer, Double>; null price or guest -> -1.0 (PriceCorrelationDataQualityUdf.java:12-14)."""

    serialVersionUID = 4949954702581973224
    name = "priceCorrelationRule"

    def call(self, price, guest):
        if price is None or guest is None:
            return -1.0
        return PriceCorrelationDataQualityService.checkPriceRange(price, guest)

    def call_columns(self, ctx, price, guest, price_valid=None, guest_valid=None, params=None):
        out = _alloc_like(price, np.float64, "float64")
        p = C.byref(params) if params is not None else None
        capi.check(capi.lib().dq4ml_udf_priceCorrelationRule(
            ctx.handle, capi.buf_ptr(price), capi.buf_ptr(guest), capi.buf_ptr(price_valid), capi.buf_ptr(guest_valid),
            capi.buf_len(price), capi.buf_ptr(out), p))
        return out


class UDFRegistration:
    """spark.udf(): register(name, udf, returnType) and lookup by name (App:46-49)."""

    def __init__(self):
        self._udfs = {}

    def register(self, name, udf, return_type=DoubleType):
        if return_type != DoubleType:
            raise ValueError("the lab registers both rules with DataTypes.DoubleType")
        self._udfs[name] = udf
        return udf

    def get(self, name):
        if name not in self._udfs:
            raise KeyError(f"Undefined function: '{name}'")  # Spark: AnalysisException
        return self._udfs[name]


def callUDF(ctx, registry, name, *cols, **kw):
    """functions.callUDF(name, cols...) on whole columns (App:69: callUDF("minimumPriceRule", price);
    App:87: callUDF("priceCorrelationRule", price, guest))."""
    return registry.get(name).call_columns(ctx, *cols, **kw)


def default_params():
    return capi.lib().dq4ml_dq_params_default()


def rule1_only_params():
    """Thresholds under which dq_filter applies rule 1 + its filter only (App:68-78)."""
    p = default_params()
    p.min_guest = -(2 ** 31)  # rule 2 can never fire
    return p


def rule2_only_params():
    """Thresholds under which dq_filter applies rule 2 + its filter only (App:86-90)."""
    p = default_params()
    p.min_price = float("-inf")  # rule 1 passes every price through
    return p


def dq_filter(ctx, guest, price, params=None):
    """The DQ section in one pass (App:68-90).  Returns (guest', price', n_after_rule1); order preserved."""
    n = capi.buf_len(guest)
    go = _alloc_like(guest, np.int32, "int32")
    po = _alloc_like(price, np.float64, "float64")
    n_out, n1 = C.c_int64(0), C.c_int64(0)
    p = C.byref(params) if params is not None else None
    capi.check(capi.lib().dq4ml_dq_filter(ctx.handle, capi.buf_ptr(guest), capi.buf_ptr(price), n, p, capi.buf_ptr(go),
                                           capi.buf_ptr(po), C.byref(n_out), C.byref(n1)))
    m = int(n_out.value)
    return go[:m], po[:m], int(n1.value)


def dq_filter_global_offset(ctx, n_out_local):
    """Row-sharded DQ section (one rank per GPU): collective; returns (offset of this rank's survivors in the global
    order-preserving output, total survivors over all ranks)."""
    off, tot = C.c_int64(0), C.c_int64(0)
    capi.check(capi.lib().dq4ml_dq_filter_global_offset(ctx.handle, int(n_out_local), C.byref(off), C.byref(tot)))
    return int(off.value), int(tot.value)
