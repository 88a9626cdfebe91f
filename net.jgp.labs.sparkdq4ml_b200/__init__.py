"""net.jgp.labs.sparkdq4ml on B200: the lab's DQ rules + LinearRegression hot path as sm_100a CUDA kernels behind a
C ABI (csrc/, include/dq4ml.h), with this thin host-side mirror of the lab's API for tests and benchmarks.

The directory keeps the reference's dotted package name, which Python cannot import by statement; the repo-root shim
`sparkdq4ml_b200.py` registers it as the importable package `sparkdq4ml_b200`.
"""
from . import _capi as capi  # noqa: F401
from .runtime import Context, shard_bounds  # noqa: F401
from .dq import (  # noqa: F401
    DoubleType, MinimumPriceDataQualityService, PriceCorrelationDataQualityService, MinimumPriceDataQualityUdf,
    PriceCorrelationDataQualityUdf, UDFRegistration, callUDF, dq_filter, dq_filter_global_offset)
from .ml import (  # noqa: F401
    Frame, VectorAssembler, LinearRegression, LinearRegressionModel, LinearRegressionTrainingSummary)
from .app import DataQuality4MachineLearningApp, parse_lab_csv  # noqa: F401
