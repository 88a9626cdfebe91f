"""GPU parity, loss = huber (LinearRegression.setLoss("huber"), the parameter SURVEY.md 8f lists after weightCol): the
summarizer + HuberAggregator passes on the GPU and the C++ L-BFGS-B against the oracle's restatement of Spark 2.4.4 /
Breeze 0.13.2 (same optimiser states, coefficients / intercept / scale / objective history within 1e-8), and against
scikit-learn's HuberRegressor, which minimises the same objective (times 2n, with alpha = n * regParam) -- that pins the
optimum independently of either restatement.  PARITY UNPINNED by the reference, like every Spark-equality claim here."""
import numpy as np
import pytest
import torch

from conftest import rel_err, scale_err
from oracle import passes, spark_ml

pytestmark = pytest.mark.gpu

TOL = 1e-8
TOL_WIDE = 1e-6


def table(n, d, seed=0):
    """Synthetic rows WITHOUT the DQ rules applied: the injected price anomalies stay in as outliers."""
    guest, price, extra = passes.synth_generate(n, d=d, p_low=0.03, p_high=0.03, seed_data=20200601 + seed)
    feats = [guest] + [np.ascontiguousarray(extra[k]) for k in range(d - 1)]
    return feats, price


def fit_gpu(pkg, ctx, feats, label, params, weight=None, dq=None):
    cols = {"label": torch.from_numpy(label).cuda()}
    names = []
    for j, f in enumerate(feats):
        cols[f"f{j}"] = torch.from_numpy(np.ascontiguousarray(f)).cuda()
        names.append(f"f{j}")
    if weight is not None:
        cols["w"] = torch.from_numpy(weight).cuda()
    df = pkg.Frame(ctx, cols)
    if dq is not None:
        df = df.with_fused_dq("f0", "label")
    df = pkg.VectorAssembler().setInputCols(names).transform(df)
    lr = pkg.LinearRegression().setLoss("huber").setEpsilon(params.get("epsilon", 1.35)).setMaxIter(params.get("max_iter", 100))
    lr.setRegParam(params.get("reg_param", 0.0)).setTol(params.get("tol", 1e-6)).setFitIntercept(params.get("fit_intercept", True))
    lr.setStandardization(params.get("standardization", True)).setSolver(params.get("solver", "auto"))
    if weight is not None:
        lr.setWeightCol("w")
    return lr.fit(df)


CASES = [
    (1, 60_000, dict()),
    (4, 80_003, dict(reg_param=0.1)),
    (4, 80_003, dict(reg_param=0.1, standardization=False)),
    (4, 50_000, dict(fit_intercept=False, epsilon=2.0)),
    (10, 30_001, dict(reg_param=0.05)),   # 9..12 columns: lane-group kernels (weighted: the weight + 8..16 features)
    (12, 20_000, dict(reg_param=0.02, fit_intercept=False)),
    (16, 120_000, dict(reg_param=0.05, solver="l-bfgs")),
    (24, 25_003, dict(reg_param=0.03)),   # 17..32 columns: groups of four lanes
    (32, 40_000, dict(reg_param=0.01)),
    (40, 30_000, dict(reg_param=0.02)),   # wider than one streaming launch: column-walking row pass + gradient groups
    (70, 20_003, dict(reg_param=0.05, fit_intercept=False)),
]


@pytest.mark.parametrize("d,n,params", CASES)
@pytest.mark.parametrize("weighted", [False, True])
def test_huber_fit_vs_oracle(pkg, ctx, parity_log, d, n, params, weighted):
    feats, price = table(n, d)
    w = None
    if weighted:
        rng = np.random.default_rng(5)
        w = rng.integers(0, 3, n).astype(np.float64) * rng.choice([0.5, 1.0, 2.25], n)
    p = spark_ml.LRParams(loss="huber", solver="l-bfgs", **{k: v for k, v in params.items() if k != "solver"})
    ref = spark_ml.linear_regression_fit(feats, price, p, weight=w)
    model = fit_gpu(pkg, ctx, feats, price, params, weight=w)
    assert model.solver_used == ref.solver_used == "lbfgsb"
    s = model.summary()
    assert s.totalIterations() == ref.total_iterations and model.data_passes == ref.data_passes
    errs = [rel_err(s.objectiveHistory(), ref.objective_history), scale_err(model.coefficients(), ref.coefficients),
            abs(model.intercept() - ref.intercept) / max(abs(ref.intercept), 1e-300) if ref.intercept != 0 else abs(model.intercept()),
            rel_err(model.scale(), ref.scale)]
    parity_log(f"huber_d{d}_n{n}_w{int(weighted)}_{params}", "objectiveHistory / coef scale-rel / intercept / scale", errs)
    # up to 32 streamed columns the sums differ from the oracle's in the last bits only (measured <= 9e-11); the two-pass
    # path of wider tables sums in another order and ~40 L-BFGS-B iterations amplify that (measured 4e-8 at d = 40)
    assert errs[0] < TOL and max(errs) < (TOL if d <= 32 else TOL_WIDE), errs


def test_huber_with_fused_dq_equals_filtered_table(pkg, ctx):
    n, d = 70_000, 4
    guest, price, extra = passes.synth_generate(n, d=d, p_low=0.05, p_high=0.05)
    feats = [guest] + [np.ascontiguousarray(extra[k]) for k in range(d - 1)]
    keep = spark_ml._dq_keep_mask((guest, price), n)
    params = dict(reg_param=0.02)
    fused = fit_gpu(pkg, ctx, feats, price, params, dq=True)
    filt = fit_gpu(pkg, ctx, [np.ascontiguousarray(f[keep]) for f in feats], np.ascontiguousarray(price[keep]), params)
    assert fused.num_rows == filt.num_rows == int(keep.sum())
    assert scale_err(fused.coefficients(), filt.coefficients()) < 1e-9 and rel_err(fused.scale(), filt.scale()) < 1e-9


def test_huber_optimum_vs_scikit_learn(pkg, ctx):
    """Same objective: sklearn minimises sum_i (sigma + H(r_i / sigma) sigma) + alpha ||w||^2 = 2n x Spark's objective
    with alpha = n regParam (standardization off, so both penalise the raw coefficients)."""
    from sklearn.linear_model import HuberRegressor
    n, d = 40_000, 3
    feats, price = table(n, d, seed=3)
    x = np.stack([np.asarray(f, dtype=np.float64) for f in feats], axis=1)
    for reg in (0.0, 0.2):
        model = fit_gpu(pkg, ctx, feats, price, dict(reg_param=reg, standardization=False, tol=1e-12, max_iter=500))
        hr = HuberRegressor(epsilon=1.35, alpha=reg * n, tol=1e-12, max_iter=20000).fit(x, price)
        np.testing.assert_allclose(model.coefficients(), hr.coef_, rtol=2e-4, atol=2e-5)
        np.testing.assert_allclose(model.intercept(), hr.intercept_, rtol=2e-4)
        np.testing.assert_allclose(model.scale(), hr.scale_, rtol=2e-4)


def test_huber_parameter_checks(pkg, ctx):
    feats, price = table(1000, 2)
    for bad in (dict(solver="normal"), dict(epsilon=1.0)):
        with pytest.raises((pkg.capi.Dq4mlError, ValueError)):
            fit_gpu(pkg, ctx, feats, price, bad)
    cols = {"label": torch.from_numpy(price).cuda(), "f0": torch.from_numpy(feats[0]).cuda()}
    df = pkg.VectorAssembler().setInputCols(["f0"]).transform(pkg.Frame(ctx, cols))
    with pytest.raises(pkg.capi.Dq4mlError):  # huber supports L2 only
        pkg.LinearRegression().setLoss("huber").setRegParam(0.1).setElasticNetParam(0.5).fit(df)
