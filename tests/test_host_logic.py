"""CPU: the C-ABI library loads, exports every symbol include/dq4ml.h declares, and its host-only entry points (scalar
rules, WLS solve = Cholesky / OWL-QN / L-BFGS over the Gram) agree with the oracle.  No GPU compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from conftest import ROOT, RTOL, rel_err
from oracle import passes, spark_ml


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "dq4ml.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = set(re.findall(r"\b(dq4ml_[A-Za-z0-9_]+)\s*\(", text))
    names.discard("dq4ml_allreduce_fn")
    return sorted(names)


def test_library_loads_and_exports_every_declared_symbol(pkg):
    L = pkg.capi.lib()
    assert L.dq4ml_abi_version() == 2
    syms = declared_symbols()
    assert len(syms) >= 40
    raw = C.CDLL(pkg.capi.LIB_PATH)
    for s in syms:
        assert hasattr(raw, s), f"{s} declared in include/dq4ml.h but not exported"
    assert set(pkg.capi.SIGNATURES) == set(syms), set(pkg.capi.SIGNATURES) ^ set(syms)


def test_no_cpu_fallback_without_device(pkg):
    """On a box without a GPU the product must fail loudly instead of computing on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.capi.Dq4mlError) as ei:
        pkg.Context(0)
    assert ei.value.status == pkg.capi.ERR_NO_DEVICE and "no CPU fallback" in str(ei.value)


@settings(max_examples=300, deadline=None)
@given(price=st.floats(allow_nan=True, allow_infinity=True), guest=st.integers(-2**31, 2**31 - 1))
def test_scalar_rules_bitwise_equal_oracle(pkg, price, guest):
    a = pkg.MinimumPriceDataQualityService.checkMinimumPrice(price)
    b = passes.check_minimum_price(price)
    assert np.float64(a).tobytes() == np.float64(b).tobytes()
    a = pkg.PriceCorrelationDataQualityService.checkPriceRange(price, guest)
    b = passes.check_price_range(price, guest)
    assert np.float64(a).tobytes() == np.float64(b).tobytes()


def test_udf_adapters_row_contract(pkg):
    u1, u2 = pkg.MinimumPriceDataQualityUdf(), pkg.PriceCorrelationDataQualityUdf()
    assert u1.call(19.0) == -1.0 and u1.call(20.0) == 20.0
    with pytest.raises(pkg.capi.NullValueError):
        u1.call(None)
    assert u2.call(None, 3) == -1.0 and u2.call(95.0, None) == -1.0 and u2.call(95.0, 13) == -1.0 and u2.call(95.0, 14) == 95.0
    assert u1.serialVersionUID == -201966159201746851 and u2.serialVersionUID == 4949954702581973224
    reg = pkg.UDFRegistration()
    reg.register("minimumPriceRule", u1, pkg.DoubleType)
    reg.register("priceCorrelationRule", u2, pkg.DoubleType)
    assert reg.get("minimumPriceRule") is u1
    with pytest.raises(KeyError):
        reg.get("noSuchRule")


def _estimator(pkg, params):
    lr = pkg.LinearRegression().setMaxIter(params.get("max_iter", 100)).setRegParam(params.get("reg_param", 0.0))
    lr.setElasticNetParam(params.get("elastic_net_param", 0.0)).setTol(params.get("tol", 1e-6))
    lr.setFitIntercept(params.get("fit_intercept", True)).setStandardization(params.get("standardization", True))
    lr.setSolver(params.get("solver", "auto"))
    return lr


@pytest.mark.parametrize("name", ["abstract", "small", "full", "synth_small"])
def test_wls_solve_matches_oracle_on_golden_moments(pkg, golden, name):
    from sparkdq4ml_b200 import ml
    doc = golden(name if name == "synth_small" else "dataset_" + name)
    if name == "synth_small":
        guest, price, extra = passes.synth_generate(doc["n"], d=doc["d"], p_low=doc["p_low"], p_high=doc["p_high"])
        feats = [guest] + [extra[k] for k in range(doc["d"] - 1)]
        mom = passes.wls_moments(feats, price, dq=(guest, price))
        d = doc["d"]
    else:
        mom, d = np.asarray(doc["wls_moments"]), 1
    for fit_name, ref in doc["fits"].items():
        if ref["params"].get("solver") == "l-bfgs":
            continue
        model = ml.wls_solve(mom, d, _estimator(pkg, ref["params"]))
        assert model.solver_used == ref["solver_used"], fit_name
        assert rel_err(model.coefficients(), ref["coefficients"]) < RTOL, fit_name
        assert rel_err(model.intercept(), ref["intercept"]) < RTOL, fit_name
        hist = model.summary if False else None
        h = np.zeros(64)
        k = pkg.capi.lib().dq4ml_linreg_model_objective_history(model._h, h.ctypes.data, 64)
        assert k == ref["total_iterations"], fit_name
        assert rel_err(h[:k], ref["objective_history"]) < RTOL, fit_name


@settings(max_examples=25, deadline=None)
@given(seed=st.integers(0, 10_000), d=st.integers(1, 6), reg=st.sampled_from([0.0, 0.05, 0.5, 2.0]),
       alpha=st.sampled_from([0.0, 0.3, 1.0]), intercept=st.booleans(), std=st.booleans())
def test_wls_solve_property(pkg, seed, d, reg, alpha, intercept, std):
    """Random small problems: host C++ solver == oracle's restatement, iterate for iterate."""
    from sparkdq4ml_b200 import ml
    rng = np.random.default_rng(seed)
    n = 200
    X = rng.normal(size=(n, d)) * rng.uniform(0.5, 5.0, size=d) + rng.uniform(-3, 3, size=d)
    y = X @ rng.normal(size=d) + 3.0 + rng.normal(size=n)
    feats = [np.ascontiguousarray(X[:, j]) for j in range(d)]
    mom = passes.wls_moments(feats, y)
    kw = dict(max_iter=50, reg_param=reg, elastic_net_param=alpha, fit_intercept=intercept, standardization=std)
    ref = spark_ml.wls_fit_from_moments(mom, d, spark_ml.LRParams(**kw))
    model = ml.wls_solve(mom, d, _estimator(pkg, kw))
    assert model.solver_used == ref.solver_used
    h = np.zeros(128)
    k = pkg.capi.lib().dq4ml_linreg_model_objective_history(model._h, h.ctypes.data, 128)
    assert k == len(ref.objective_history)
    assert rel_err(h[:k], ref.objective_history) < 1e-8
    np.testing.assert_allclose(model.coefficients(), ref.coefficients, rtol=1e-7, atol=1e-10)
    np.testing.assert_allclose(model.intercept(), ref.intercept, rtol=1e-7, atol=1e-10)


def test_wls_solve_special_cases(pkg):
    from sparkdq4ml_b200 import ml
    lr = pkg.LinearRegression().setRegParam(1.0).setElasticNetParam(1.0)
    # constant label -> zero coefficients, intercept = mean, history [0.0]
    x = np.arange(10, dtype=np.float64)
    y = np.full(10, 7.5)
    m = ml.wls_solve(passes.wls_moments([x], y), 1, lr)
    assert m.coefficients().tolist() == [0.0] and m.intercept() == 7.5 and m.solver_used == "trivial"
    # empty training set -> error like Spark's "Training dataset is empty"
    with pytest.raises(pkg.capi.Dq4mlError) as ei:
        ml.wls_solve(np.zeros(8), 1, lr)
    assert ei.value.status == pkg.capi.ERR_EMPTY
    # constant feature -> its coefficient is 0
    xc = np.full(10, 3.0)
    m = ml.wls_solve(passes.wls_moments([x, xc], 2 * x + 1), 2, pkg.LinearRegression())
    ref = spark_ml.wls_fit_from_moments(passes.wls_moments([x, xc], 2 * x + 1), 2, spark_ml.LRParams())
    assert m.solver_used == ref.solver_used
    np.testing.assert_allclose(m.coefficients(), ref.coefficients, rtol=1e-9, atol=1e-12)
    # parameter validation
    with pytest.raises(ValueError):
        pkg.LinearRegression().setElasticNetParam(1.5)
    with pytest.raises(ValueError):
        pkg.LinearRegression().setSolver("sgd")


def test_predict_and_model_create(pkg):
    m = pkg.LinearRegressionModel.from_coefficients([4.878439749218744], 23.963255452175247)
    assert m.predict([40.0]) == pytest.approx(219.100845420925, rel=1e-13)
    assert m.numFeatures() == 1
    with pytest.raises(ValueError):
        m.predict([1.0, 2.0])
    with pytest.raises(RuntimeError):
        m.summary()


def test_shard_bounds_cover_rows_once(pkg):
    for n in (0, 1, 1023, 1024, 10_000_019):
        for world in (1, 2, 3, 8):
            cuts = [pkg.shard_bounds(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            for (a, b), (c, d) in zip(cuts, cuts[1:]):
                assert b == c and a <= b
            assert all(lo % 1024 == 0 for lo, _ in cuts)


def test_lab_csv_parser_matches_oracle(pkg, lab_csv):
    for name in ("abstract", "small", "full"):
        g, p = pkg.parse_lab_csv(lab_csv(name))
        go, po = passes.csv_parse(lab_csv(name))
        assert g.tolist() == go.tolist() and p.tobytes() == po.tobytes()


def test_inference_statistics_match_numpy_scipy(pkg):
    """coefficientStandardErrors / tValues / pValues (host-only entry point) against the textbook OLS formulas."""
    from scipy import stats
    from sparkdq4ml_b200 import ml
    rng = np.random.default_rng(5)
    n, d = 500, 5
    X = rng.normal(size=(n, d)) * rng.uniform(0.5, 4.0, size=d) + rng.uniform(-2, 2, size=d)
    y = X @ np.array([1.5, 0.0, -2.0, 0.02, 0.7]) + 4.0 + rng.normal(size=n) * 1.3
    feats = [np.ascontiguousarray(X[:, j]) for j in range(d)]
    mom = passes.wls_moments(feats, y)
    model = ml.wls_solve(mom, d, pkg.LinearRegression())  # regParam 0 -> Cholesky
    assert model.solver_used == "wls-cholesky"
    Xa = np.hstack([X, np.ones((n, 1))])
    beta = np.linalg.solve(Xa.T @ Xa, Xa.T @ y)
    np.testing.assert_allclose(np.append(model.coefficients(), model.intercept()), beta, rtol=1e-9, atol=1e-11)
    res = y - Xa @ beta
    dof = n - d - 1
    se_ref = np.sqrt(np.diag(np.linalg.inv(Xa.T @ Xa)) * (res @ res) / dof)
    t_ref = beta / se_ref
    p_ref = 2.0 * stats.t.sf(np.abs(t_ref), dof)
    se, tv, pv = (np.zeros(d + 1) for _ in range(3))
    k = pkg.capi.lib().dq4ml_linreg_model_inference(model._h, float(n), float(res @ res), 1, se.ctypes.data, tv.ctypes.data,
                                                    pv.ctypes.data, d + 1)
    assert k == d + 1
    np.testing.assert_allclose(se, se_ref, rtol=1e-8)
    np.testing.assert_allclose(tv, t_ref, rtol=1e-8)
    np.testing.assert_allclose(pv, p_ref, rtol=1e-7, atol=1e-300)
    # models fitted another way have no standard errors, like Spark
    lasso = ml.wls_solve(mom, d, pkg.LinearRegression().setRegParam(0.1).setElasticNetParam(1.0))
    assert pkg.capi.lib().dq4ml_linreg_model_inference(lasso._h, float(n), 1.0, 1, None, None, None, d + 1) == pkg.capi.ERR_UNSUPPORTED
