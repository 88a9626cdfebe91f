"""GPU parity for LinearRegression.setWeightCol: weighted Gram moments, weighted summarizer / loss-gradient
passes and whole fits against the numpy restatement in oracle/spark_ml.py, plus the weighted part of the summary."""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import passes, spark_ml

pytestmark = pytest.mark.gpu


def to_dev(*arrs):
    return [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in arrs]


def table(n, d, seed=3):
    guest, price, extra = passes.synth_generate(n, d=d, p_low=0.05, p_high=0.05)
    feats = [guest] + [np.ascontiguousarray(extra[k]) for k in range(d - 1)]
    rng = np.random.default_rng(seed)
    w = rng.integers(0, 4, n).astype(np.float64) * rng.choice([0.5, 1.0, 1.75], n)  # a fifth of the rows get weight 0
    return guest, price, feats, w


def frame(pkg, ctx, feats, price, w, fused):
    dfeats, dprice, dw = to_dev(*feats), to_dev(price)[0], to_dev(w)[0]
    cols = {"guest": dfeats[0], "label": dprice, "w": dw}
    names = ["guest"]
    for k in range(1, len(feats)):
        cols[f"x{k}"] = dfeats[k]
        names.append(f"x{k}")
    df = pkg.Frame(ctx, cols)
    if fused:
        df = df.with_fused_dq("guest", "label")
    return pkg.VectorAssembler().setInputCols(names).transform(df)


def oracle_params(**kw):
    base = dict(max_iter=100, reg_param=0.0, elastic_net_param=0.0, tol=1e-6, fit_intercept=True, standardization=True,
                solver="auto")
    base.update(kw)
    return spark_ml.LRParams(**base)


def estimator(pkg, p):
    lr = pkg.LinearRegression().setMaxIter(p.max_iter).setRegParam(p.reg_param).setElasticNetParam(p.elastic_net_param)
    return lr.setTol(p.tol).setFitIntercept(p.fit_intercept).setStandardization(p.standardization).setSolver(p.solver)


@pytest.mark.parametrize("d,n", [(1, 1), (2, 31), (2, 40_003), (8, 4097), (17, 10_000), (64, 20_001), (70, 9_000),
                                 (130, 5_003)])
@pytest.mark.parametrize("fused", [False, True])
def test_weighted_normal_equation_fit_vs_oracle(pkg, ctx, d, n, fused):
    guest, price, feats, w = table(n, d)
    p = oracle_params(solver="normal", reg_param=0.0 if n < 100 else 0.05)
    if n == 1:
        w[:] = 2.0
        price[:] = 50.0
    dq = (guest, price) if fused else None
    ref = spark_ml.linear_regression_fit(feats, price, p, dq=dq, weight=w)
    model = estimator(pkg, p).setWeightCol("w").fit(frame(pkg, ctx, feats, price, w, fused))
    scale = max(1.0, float(np.max(np.abs(ref.coefficients)))) if d else 1.0
    assert float(np.max(np.abs(model.coefficients() - ref.coefficients))) / scale < 1e-7
    assert abs(model.intercept() - ref.intercept) / max(1.0, abs(ref.intercept)) < 1e-7
    keep = spark_ml._dq_keep_mask(dq, n)
    assert model.num_rows == int(keep.sum())  # WeightedLeastSquares counts zero-weight rows too


def test_weighted_fit_equals_replicated_rows(pkg, ctx):
    """Integer weights == the same rows repeated (exact for the normal-equation branch: population moments)."""
    n, d = 6_000, 5
    guest, price, feats, _ = table(n, d)
    w = np.random.default_rng(9).integers(0, 4, n).astype(np.float64)
    idx = np.repeat(np.arange(n), w.astype(np.int64))
    p = oracle_params(solver="normal", reg_param=0.3, elastic_net_param=0.5)
    lr = estimator(pkg, p)
    mw = lr.setWeightCol("w").fit(frame(pkg, ctx, feats, price, w, True))
    rep = [np.ascontiguousarray(f[idx]) for f in feats]
    mr = estimator(pkg, p).fit(frame(pkg, ctx, rep, np.ascontiguousarray(price[idx]), np.ones(idx.size), True))
    np.testing.assert_allclose(mw.coefficients(), mr.coefficients(), rtol=1e-8, atol=1e-10)
    assert abs(mw.intercept() - mr.intercept()) < 1e-7 * max(1.0, abs(mr.intercept()))


# 8..16 features: lane-group kernels with the weight as a row scalar (8 and 16: the weighted summarizer as well)
@pytest.mark.parametrize("d,n", [(1, 2_000), (3, 50_001), (8, 9_001), (9, 4_096), (12, 7_000), (15, 8_000), (16, 7_001), (31, 6_001), (32, 6_001), (40, 9_000),
                                 (70, 5_003)])
@pytest.mark.parametrize("reg", [(0.0, 0.0), (0.2, 0.0), (0.3, 0.8)])
@pytest.mark.parametrize("fused", [False, True])
def test_weighted_lbfgs_fit_vs_oracle(pkg, ctx, d, n, reg, fused):
    guest, price, feats, w = table(n, d)
    p = oracle_params(solver="l-bfgs", reg_param=reg[0], elastic_net_param=reg[1], tol=1e-9, max_iter=60)
    dq = (guest, price) if fused else None
    ref = spark_ml.linear_regression_fit(feats, price, p, dq=dq, weight=w)
    model = estimator(pkg, p).setWeightCol("w").fit(frame(pkg, ctx, feats, price, w, fused))
    keep = spark_ml._dq_keep_mask(dq, n)
    assert model.num_rows == int((keep & (w != 0)).sum())  # the summarizer skips zero-weight rows
    # both sides run the same optimiser on sums that differ in the last bits: compare at the optimiser's tolerance
    scale = max(1.0, float(np.max(np.abs(ref.coefficients))))
    assert float(np.max(np.abs(model.coefficients() - ref.coefficients))) / scale < 2e-5
    assert abs(model.intercept() - ref.intercept) / max(1.0, abs(ref.intercept)) < 2e-5
    hist = model.summary().objectiveHistory()
    assert abs(hist[0] - ref.objective_history[0]) < 1e-9 * max(1.0, abs(ref.objective_history[0]))


def test_weighted_lbfgs_wide_table_nan_weight_is_an_error(pkg, ctx):
    """Tables of 32 and more feature columns take the grouped weighted summarizer and the two-pass weighted
    loss/gradient (the weight no longer fits as column 0 of one streaming launch): same require(w >= 0) behaviour."""
    guest, price, feats, w = table(2_000, 40)
    w[77] = -0.5
    with pytest.raises(pkg.capi.Dq4mlError) as ei:
        pkg.LinearRegression().setSolver("l-bfgs").setWeightCol("w").fit(frame(pkg, ctx, feats, price, w, False))
    assert ei.value.status == pkg.capi.ERR_INVALID


@pytest.mark.parametrize("solver", ["normal", "l-bfgs"])
def test_negative_or_nan_weight_is_an_error(pkg, ctx, solver):
    guest, price, feats, w = table(3_000, 3)
    for bad in (-1.0, float("nan")):
        w2 = w.copy()
        w2[1234] = bad
        with pytest.raises(pkg.capi.Dq4mlError) as ei:
            pkg.LinearRegression().setSolver(solver).setWeightCol("w").fit(frame(pkg, ctx, feats, price, w2, False))
        assert ei.value.status == pkg.capi.ERR_INVALID


def test_all_weights_one_equals_unweighted_bitwise(pkg, ctx):
    """The Gram kernel with w = 1 must give the moments of the unweighted kernel exactly (sqrt(1) scaling is exact)."""
    guest, price, feats, _ = table(30_011, 20)
    ones = np.ones(price.size)
    p = oracle_params(solver="normal")
    a = estimator(pkg, p).setWeightCol("w").fit(frame(pkg, ctx, feats, price, ones, True))
    b = estimator(pkg, p).fit(frame(pkg, ctx, feats, price, ones, True))
    assert a.coefficients().tobytes() == b.coefficients().tobytes() and a.intercept() == b.intercept()


@pytest.mark.parametrize("d", [3, 40])
def test_weighted_summary_surface_vs_numpy(pkg, ctx, d):
    """Spark 2.4.4 keeps RegressionMetrics unweighted; coefficientStandardErrors use rss = sum w r^2 and
    devianceResiduals the extremes of sqrt(w) r."""
    n = 40_000
    guest, price, feats, w = table(n, d)
    keep = spark_ml._dq_keep_mask((guest, price), n)
    X = np.stack([np.asarray(f, dtype=np.float64)[keep] for f in feats], axis=1)
    y, wk = price[keep], w[keep]
    model = pkg.LinearRegression().setSolver("normal").setWeightCol("w").fit(frame(pkg, ctx, feats, price, w, True))
    s = model.summary()
    m = len(y)
    Xa = np.hstack([X, np.ones((m, 1))])
    sw = np.sqrt(wk)
    beta = np.linalg.lstsq(Xa * sw[:, None], y * sw, rcond=None)[0]
    np.testing.assert_allclose(np.append(model.coefficients(), model.intercept()), beta, rtol=1e-7, atol=1e-9)
    res = y - Xa @ beta
    assert s.numInstances() == m and s.degreesOfFreedom() == m - d - 1
    assert rel_err(s.meanSquaredError(), (res @ res) / m) < 1e-8  # unweighted, as in Spark 2.4.4
    np.testing.assert_allclose(s.devianceResiduals(), [(sw * res).min(), (sw * res).max()], rtol=1e-6, atol=1e-7)
    dof = m - d - 1
    atwa = (Xa * wk[:, None]).T @ Xa
    se_ref = np.sqrt(np.diag(np.linalg.inv(atwa)) * float((wk * res * res).sum()) / dof)
    np.testing.assert_allclose(s.coefficientStandardErrors(), se_ref, rtol=1e-6)
    np.testing.assert_allclose(s.tValues(), beta / se_ref, rtol=1e-6)


def test_weighted_residual_stats_host_buffers(pkg, ctx):
    rng = np.random.default_rng(5)
    n = 10_001
    r, w = rng.normal(size=n), rng.uniform(0, 3, n)
    out = np.zeros(3)
    pkg.capi.check(pkg.capi.lib().dq4ml_weighted_residual_stats(ctx.handle, r.ctypes.data, w.ctypes.data, n, None,
                                                                out.ctypes.data))
    s = np.sqrt(w) * r
    assert rel_err(out[0], float((w * r * r).sum())) < 1e-12
    assert out[1] == s.min() and out[2] == s.max()
