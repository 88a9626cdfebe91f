"""Host-side mirror of the Spark ML surface the lab uses (DataQuality4MachineLearningApp.java:101-154), over the C ABI.

  Frame                          ~ Dataset<Row> restricted to what the path needs: named columns + optional fused DQ
  VectorAssembler                ~ org.apache.spark.ml.feature.VectorAssembler            (App:106-113)
  LinearRegression               ~ org.apache.spark.ml.regression.LinearRegression        (App:120-126)
  LinearRegressionModel          ~ ...LinearRegressionModel                               (App:126-129,141-151)
  LinearRegressionTrainingSummary~ ...LinearRegressionTrainingSummary                     (App:132-139)

Columns are numpy arrays (host) or torch tensors (host or cuda:<ctx.device>); nothing is copied or re-laid-out: the
"features" vector column is a table of column references.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi as capi


class Frame:
    """Named columns of equal length.  `dq` marks the frame as "DQ rules pending": consumers fuse the two rules and
    their filters into their own pass instead of reading a filtered copy."""

    def __init__(self, ctx, columns: dict, dq=None, dq_params=None):
        self.ctx = ctx
        self.columns = dict(columns)
        self.dq = dq  # (guest_col_name, price_col_name) or None
        self.dq_params = dq_params
        lens = {capi.buf_len(c) for c in self.columns.values() if not isinstance(c, FeatureTable)}
        if len(lens) > 1:
            raise ValueError(f"columns of different lengths: {lens}")
        self.n = lens.pop() if lens else 0
        self._marshalled = {}  # C descriptors of this frame's columns, built on first use (LinearRegression.fit)

    def col(self, name):
        return self.columns[name]

    def withColumn(self, name, col):
        cols = dict(self.columns)
        cols[name] = col
        return Frame(self.ctx, cols, self.dq, self.dq_params)

    def withColumnRenamed(self, old, new):
        cols = {(new if k == old else k): v for k, v in self.columns.items()}
        return Frame(self.ctx, cols, self.dq, self.dq_params)

    def with_fused_dq(self, guest="guest", price="price", params=None):
        return Frame(self.ctx, self.columns, (guest, price), params)

    def count(self):
        return self.n

    def _dq_struct(self):
        if self.dq is None:
            return None, ()
        return capi.make_dq(self.columns[self.dq[0]], self.columns[self.dq[1]], self.dq_params)


class FeatureTable:
    """What VectorAssembler produces here: references to the input columns, in order (never a packed copy)."""

    def __init__(self, cols, names):
        self.cols = list(cols)
        self.names = list(names)

    def __len__(self):
        return len(self.cols)


class VectorAssembler:
    def __init__(self):
        self._input_cols = []
        self._output_col = "features"

    def setInputCols(self, cols):
        self._input_cols = list(cols)
        return self

    def setOutputCol(self, name):
        self._output_col = name
        return self

    def getInputCols(self):
        return list(self._input_cols)

    def getOutputCol(self):
        return self._output_col

    def transform(self, df: Frame) -> Frame:
        for c in self._input_cols:
            if c not in df.columns:
                raise KeyError(f"VectorAssembler: input column '{c}' does not exist")
        table = FeatureTable([df.columns[c] for c in self._input_cols], self._input_cols)
        return df.withColumn(self._output_col, table)


class LinearRegression:
    """Same setters and defaults as Spark 2.4.4 (maxIter 100, regParam 0, elasticNetParam 0, tol 1e-6, fitIntercept,
    standardization, solver auto, featuresCol "features", labelCol "label", predictionCol "prediction")."""

    def __init__(self):
        self._max_iter, self._reg_param, self._elastic_net, self._tol = 100, 0.0, 0.0, 1e-6
        self._fit_intercept, self._standardization, self._solver = True, True, "auto"
        self._features_col, self._label_col, self._prediction_col = "features", "label", "prediction"
        self._weight_col = None
        self._loss, self._epsilon = "squaredError", 1.35

    def setLoss(self, v):
        """"squaredError" (default) or "huber" (Spark 2.3+: robust regression, l-bfgs only, L2 regularisation only)."""
        if v not in capi.LOSSES:
            raise ValueError(f"loss must be one of {sorted(capi.LOSSES)}")
        self._loss = v
        return self

    def getLoss(self):
        return self._loss

    def setEpsilon(self, v):
        if not float(v) > 1.0:
            raise ValueError("epsilon must be > 1.0")
        self._epsilon = float(v)
        return self

    def getEpsilon(self):
        return self._epsilon

    def setWeightCol(self, v):
        """Instance weights (Spark: unset / empty = every weight 1.0; the lab never sets it)."""
        self._weight_col = v if v else None
        return self

    def getWeightCol(self):
        return self._weight_col or ""

    def setMaxIter(self, v):
        self._max_iter = int(v)
        return self

    def setRegParam(self, v):
        self._reg_param = float(v)
        return self

    def setElasticNetParam(self, v):
        if not 0.0 <= float(v) <= 1.0:
            raise ValueError("elasticNetParam must be in [0, 1]")
        self._elastic_net = float(v)
        return self

    def setTol(self, v):
        self._tol = float(v)
        return self

    def setFitIntercept(self, v):
        self._fit_intercept = bool(v)
        return self

    def setStandardization(self, v):
        self._standardization = bool(v)
        return self

    def setSolver(self, v):
        if v not in capi.SOLVERS:
            raise ValueError(f"solver must be one of {sorted(capi.SOLVERS)}")
        self._solver = v
        return self

    def setFeaturesCol(self, v):
        self._features_col = v
        return self

    def setLabelCol(self, v):
        self._label_col = v
        return self

    def setPredictionCol(self, v):
        self._prediction_col = v
        return self

    def getMaxIter(self):
        return self._max_iter

    def getRegParam(self):
        return self._reg_param

    def getElasticNetParam(self):
        return self._elastic_net

    def getTol(self):
        return self._tol

    def _c_params(self):
        return capi.LrParams(self._max_iter, self._reg_param, self._elastic_net, self._tol, int(self._fit_intercept),
                             int(self._standardization), capi.SOLVERS[self._solver], capi.LOSSES[self._loss],
                             self._epsilon)

    def fit(self, df: Frame) -> "LinearRegressionModel":
        feats = df.columns[self._features_col]
        if not isinstance(feats, FeatureTable):
            raise TypeError(f"column '{self._features_col}' must come from VectorAssembler")
        # the C descriptors of a frame's columns are built once per (frame, column names): a Frame never changes, and
        # on a 0.2 ms fit the ctypes marshalling would otherwise be a visible share of the step
        key = (self._features_col, self._label_col, self._weight_col)
        cached = df._marshalled.get(key)
        if cached is None:
            label = df.columns[self._label_col]
            tab, keep = capi.make_columns(feats.cols)
            dqs, keep2 = df._dq_struct()
            weight = df.columns[self._weight_col] if self._weight_col else None
            cached = (C.byref(tab), capi.buf_ptr(label), capi.buf_ptr(weight), C.byref(dqs) if dqs is not None else None,
                      (tab, keep, dqs, keep2, label, weight))
            df._marshalled[key] = cached
        tab_ref, label_ptr, weight_ptr, dq_ref, _ = cached
        params = self._c_params()
        handle = C.c_void_p()
        capi.check(capi.lib().dq4ml_linreg_fit_weighted(df.ctx.handle, tab_ref, label_ptr, weight_ptr, df.n,
                                                         C.byref(params), dq_ref, C.byref(handle)))
        model = LinearRegressionModel(handle, self)
        model._training = df
        return model


class LinearRegressionModel:
    def __init__(self, handle, estimator: LinearRegression):
        self._h = handle
        self._est = estimator
        self._training = None
        self._summary = None

    def __del__(self):
        try:
            if self._h:
                capi.lib().dq4ml_linreg_model_destroy(self._h)
                self._h = None
        except Exception:
            pass

    @classmethod
    def from_coefficients(cls, coefficients, intercept, estimator=None):
        c = np.ascontiguousarray(coefficients, dtype=np.float64)
        handle = C.c_void_p()
        capi.check(capi.lib().dq4ml_linreg_model_create(c.ctypes.data, c.size, float(intercept), C.byref(handle)))
        return cls(handle, estimator or LinearRegression())

    def numFeatures(self):
        return int(capi.lib().dq4ml_linreg_model_num_features(self._h))

    def coefficients(self):
        c = np.zeros(self.numFeatures(), dtype=np.float64)
        capi.check(capi.lib().dq4ml_linreg_model_coefficients(self._h, c.ctypes.data))
        return c

    def intercept(self):
        return float(capi.lib().dq4ml_linreg_model_intercept(self._h))

    def scale(self):
        """LinearRegressionModel.scale: the huber loss's sigma (1.0 for squaredError)."""
        return float(capi.lib().dq4ml_linreg_model_scale(self._h))

    def getRegParam(self):
        return self._est.getRegParam()

    def getTol(self):
        return self._est.getTol()

    def getElasticNetParam(self):
        return self._est.getElasticNetParam()

    def getMaxIter(self):
        return self._est.getMaxIter()

    @property
    def solver_used(self):
        return capi.lib().dq4ml_linreg_model_solver_used(self._h).decode()

    @property
    def data_passes(self):
        return int(capi.lib().dq4ml_linreg_model_data_passes(self._h))

    @property
    def num_rows(self):
        return int(capi.lib().dq4ml_linreg_model_num_rows(self._h))

    def predict(self, features):
        """model.predict(Vectors.dense(40.0)) (App:150-151)."""
        x = np.ascontiguousarray(np.atleast_1d(np.asarray(features, dtype=np.float64)))
        if x.size != self.numFeatures():
            raise ValueError(f"expected {self.numFeatures()} features, got {x.size}")
        return float(capi.lib().dq4ml_linreg_predict1(self._h, x.ctypes.data))

    def transform(self, df: Frame) -> Frame:
        """Adds the prediction column (App:129)."""
        feats = df.columns[self._est._features_col]
        tab, keep = capi.make_columns(feats.cols)
        ref = feats.cols[0] if len(feats.cols) else df.columns[self._est._label_col]
        if capi.is_torch(ref):
            import torch
            pred = torch.empty(df.n, dtype=torch.float64, device=ref.device)
        else:
            pred = np.empty(df.n, dtype=np.float64)
        capi.check(capi.lib().dq4ml_linreg_transform(df.ctx.handle, self._h, C.byref(tab), df.n, capi.buf_ptr(pred)))
        return df.withColumn(self._est._prediction_col, pred)

    def evaluate(self, df: Frame, want_residuals=False):
        return LinearRegressionSummary(self, df, want_residuals)

    def summary(self) -> "LinearRegressionTrainingSummary":
        if self._training is None:
            raise RuntimeError("No training summary available for this LinearRegressionModel")  # Spark: SparkException
        if self._summary is None:
            self._summary = LinearRegressionTrainingSummary(self, self._training)
        return self._summary


class LinearRegressionSummary:
    def __init__(self, model: LinearRegressionModel, df: Frame, want_residuals=True):
        self._model, self._df = model, df
        self._metrics = None
        self._residuals = None
        self._weighted = None
        self._want_residuals = want_residuals

    def _run(self, with_residuals):
        df, model = self._df, self._model
        feats = df.columns[model._est._features_col]
        label = df.columns[model._est._label_col]
        tab, keep = capi.make_columns(feats.cols)
        dqs, keep2 = df._dq_struct()
        res = None
        if with_residuals:
            if capi.is_torch(label):
                import torch
                res = torch.empty(df.n, dtype=torch.float64, device=label.device)
            else:
                res = np.empty(df.n, dtype=np.float64)
        met = np.zeros(16, dtype=np.float64)  # DQ4ML_METRICS_LEN
        capi.check(capi.lib().dq4ml_linreg_evaluate(df.ctx.handle, model._h, C.byref(tab), capi.buf_ptr(label), df.n,
                                                     C.byref(dqs) if dqs is not None else None,
                                                     int(model._est._fit_intercept), met.ctypes.data, capi.buf_ptr(res)))
        self._metrics = met
        if with_residuals:
            self._residuals = res

    def _w(self):
        """[sum w r^2, min sqrt(w) r, max sqrt(w) r] when the estimator has a weight column, else None.  Spark 2.4.4
        keeps RegressionMetrics (rmse, r2, ...) unweighted; only the standard errors and devianceResiduals use weights."""
        wc = self._model._est._weight_col
        if not wc:
            return None
        if self._weighted is None:
            df = self._df
            res = self.residuals()
            dqs, keep2 = df._dq_struct()
            out = np.zeros(3, dtype=np.float64)
            capi.check(capi.lib().dq4ml_weighted_residual_stats(df.ctx.handle, capi.buf_ptr(res),
                                                                 capi.buf_ptr(df.columns[wc]), df.n,
                                                                 C.byref(dqs) if dqs is not None else None,
                                                                 out.ctypes.data))
            self._weighted = out
        return self._weighted

    def _m(self):
        if self._metrics is None:
            self._run(self._want_residuals)
        return self._metrics

    def residuals(self):
        """label - prediction per row (App:137); rows skipped by fused DQ hold 0."""
        if self._residuals is None:
            self._run(True)
        return self._residuals

    def numInstances(self):
        return int(self._m()[0])

    def rootMeanSquaredError(self):
        return float(self._m()[5])

    def r2(self):
        return float(self._m()[6])

    def meanSquaredError(self):
        return float(self._m()[7])

    def meanAbsoluteError(self):
        return float(self._m()[8])

    def explainedVariance(self):
        return float(self._m()[9])

    def degreesOfFreedom(self):
        return int(self._m()[10])

    def r2adj(self):
        return float(self._m()[11])

    def _inference(self):
        m = self._m()
        w = self._w()
        rss = float(m[1]) if w is None else float(w[0])
        k = self._model.numFeatures() + (1 if self._model._est._fit_intercept else 0)
        se, tv, pv = (np.zeros(k, dtype=np.float64) for _ in range(3))
        capi.check(capi.lib().dq4ml_linreg_model_inference(self._model._h, float(m[0]), rss,
                                                            int(self._model._est._fit_intercept), se.ctypes.data,
                                                            tv.ctypes.data, pv.ctypes.data, k))
        return se, tv, pv

    def coefficientStandardErrors(self):
        """Standard errors of the coefficients (intercept last); normal-equation Cholesky fits only, like Spark."""
        return self._inference()[0]

    def tValues(self):
        return self._inference()[1]

    def pValues(self):
        return self._inference()[2]

    def devianceResiduals(self):
        """[min, max] of sqrt(weight) * residual, as Spark's LinearRegressionSummary.devianceResiduals."""
        w = self._w()
        if w is not None:
            return [float(w[1]), float(w[2])] if self.numInstances() > 0 else [0.0, 0.0]
        return [float(self._m()[12]), float(self._m()[13])]


class LinearRegressionTrainingSummary(LinearRegressionSummary):
    def __init__(self, model, df):
        super().__init__(model, df, want_residuals=False)

    def objectiveHistory(self):
        h = np.zeros(max(self.totalIterations(), 1), dtype=np.float64)
        n = capi.lib().dq4ml_linreg_model_objective_history(self._model._h, h.ctypes.data, h.size)
        return h[:n]

    def totalIterations(self):
        """Spark 2.4.4: objectiveHistory.length (counts the initial state)."""
        return int(capi.lib().dq4ml_linreg_model_total_iterations(self._model._h))


# ---- lower-level passes, exposed for tests / benchmarks ----------------------------------------------------------------

def wls_moments(ctx, features, label, dq=None, dq_params=None):
    tab, keep = capi.make_columns(features)
    dqs, keep2 = capi.make_dq(dq[0], dq[1], dq_params) if dq is not None else (None, ())
    out = np.zeros(int(capi.lib().dq4ml_wls_moments_len(len(features))), dtype=np.float64)
    capi.check(capi.lib().dq4ml_wls_moments(ctx.handle, C.byref(tab), capi.buf_ptr(label), capi.buf_len(label),
                                             C.byref(dqs) if dqs is not None else None, out.ctypes.data))
    return out


def wls_solve(moments, d, estimator: LinearRegression):
    """[host-only] moments -> model (WeightedLeastSquares.fit after the aggregation)."""
    mom = np.ascontiguousarray(moments, dtype=np.float64)
    params = estimator._c_params()
    handle = C.c_void_p()
    capi.check(capi.lib().dq4ml_wls_solve(mom.ctypes.data, int(d), C.byref(params), C.byref(handle)))
    return LinearRegressionModel(handle, estimator)


def summarize(ctx, features, label, dq=None, dq_params=None):
    d = len(features)
    tab, keep = capi.make_columns(features)
    dqs, keep2 = capi.make_dq(dq[0], dq[1], dq_params) if dq is not None else (None, ())
    out = np.zeros(int(capi.lib().dq4ml_summary_len(d)), dtype=np.float64)
    capi.check(capi.lib().dq4ml_summarize(ctx.handle, C.byref(tab), capi.buf_ptr(label), capi.buf_len(label),
                                           C.byref(dqs) if dqs is not None else None, out.ctypes.data))
    s = out[3:].reshape(d + 1, 7)
    return {"count": int(out[0]), "mean": s[:, 0].copy(), "variance": s[:, 1].copy(), "min": s[:, 2].copy(),
            "max": s[:, 3].copy(), "nnz": s[:, 4].copy(), "normL1": s[:, 5].copy(), "normL2": s[:, 6].copy()}


def lsq_loss_grad(ctx, features, label, coef, fstd, fmean, ystd, ymean, fit_intercept=True, dq=None, dq_params=None):
    d = len(features)
    tab, keep = capi.make_columns(features)
    dqs, keep2 = capi.make_dq(dq[0], dq[1], dq_params) if dq is not None else (None, ())
    coef = np.ascontiguousarray(coef, dtype=np.float64)
    fstd = np.ascontiguousarray(fstd, dtype=np.float64)
    fmean = np.ascontiguousarray(fmean, dtype=np.float64)
    out = np.zeros(2 + d, dtype=np.float64)
    capi.check(capi.lib().dq4ml_lsq_loss_grad(ctx.handle, C.byref(tab), capi.buf_ptr(label), capi.buf_len(label),
                                               C.byref(dqs) if dqs is not None else None, coef.ctypes.data,
                                               fstd.ctypes.data, fmean.ctypes.data, float(ystd), float(ymean),
                                               int(bool(fit_intercept)), out.ctypes.data))
    return out[0], out[1], out[2:].copy()
