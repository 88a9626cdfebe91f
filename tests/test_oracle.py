"""CPU: the oracle against the reference's fixtures, the surveyor's known-answer values, closed forms and scikit-learn."""
import hashlib

import numpy as np
import pytest

from conftest import RTOL, rel_err
from oracle import passes, spark_ml

NAMES = ["abstract", "small", "full"]


def lab_params(**kw):
    base = dict(max_iter=40, reg_param=1.0, elastic_net_param=1.0)
    base.update(kw)
    return spark_ml.LRParams(**base)


@pytest.mark.parametrize("name", NAMES)
def test_csv_and_dq_against_survey_kats(name, golden, lab_csv):
    kat = golden("survey_kats")[name]
    g, p = passes.csv_parse(lab_csv(name))
    assert len(g) == kat["n"]
    go, po, n1 = passes.dq_filter(g, p)
    assert (n1, len(go)) == (kat["n_after_rule1"], kat["n_after_rule2"])
    packed = np.zeros(len(go), dtype=[("g", "<i4"), ("p", "<f8")])
    packed["g"], packed["p"] = go, po
    assert hashlib.sha256(packed.tobytes()).hexdigest() == kat["sha256"]
    # which records each rule dropped
    r1 = passes.udf_minimum_price(p)
    drop1 = [i for i in range(len(p)) if not (r1[i] > 0)]
    assert drop1 == kat["dropped_rule1"]
    keep1 = [i for i in range(len(p)) if r1[i] > 0]
    r2 = passes.udf_price_correlation(p[keep1], g[keep1])
    drop2 = [keep1[i] for i in range(len(keep1)) if not (r2[i] > 0)]
    assert drop2 == kat["dropped_rule2"]


@pytest.mark.parametrize("name", NAMES)
def test_moments_and_lab_fit_against_survey_kats(name, golden):
    kat, doc = golden("survey_kats")[name], golden("dataset_" + name)
    go = np.asarray(doc["survivor_guest"], dtype=np.int32)
    po = np.asarray(doc["survivor_price"], dtype=np.float64)
    mom = passes.wls_moments([go], po)
    m = kat["moments"]
    assert mom[0] == kat["n_after_rule2"]
    assert rel_err(mom[3:8], [m["sum_y"], m["sum_yy"], m["sum_x"], m["sum_xy"], m["sum_xx"]]) < 1e-13
    model = spark_ml.linear_regression_fit([go], po, lab_params())
    w = kat["wls_lasso"]
    assert rel_err(model.coefficients[0], w["coefficient"]) < RTOL
    assert rel_err(model.intercept, w["intercept"]) < RTOL
    assert rel_err(model.coefficients[0] * 40 + model.intercept, w["predict_40"]) < RTOL
    assert model.total_iterations == w["n_states"]
    assert model.objective_history[0] == pytest.approx(w["objective_first"], rel=1e-12)
    assert rel_err(model.objective_history[-1], w["objective_last"]) < RTOL
    _, met = passes.predict_metrics([go], po, model.coefficients, model.intercept)
    rm = spark_ml.regression_metrics(met)
    assert rel_err(rm["rmse"], w["rmse"]) < RTOL and rel_err(rm["r2"], w["r2"]) < RTOL
    if "lbfgs_lasso" in kat:
        m2 = spark_ml.linear_regression_fit([go], po, lab_params(solver="l-bfgs"))
        assert rel_err(m2.coefficients[0], kat["lbfgs_lasso"]["coefficient"]) < RTOL
        assert rel_err(m2.intercept, kat["lbfgs_lasso"]["intercept"]) < RTOL
        n = len(po)
        assert m2.objective_history[0] == pytest.approx(0.5 * (n - 1) / n, rel=1e-12)
    if "ols" in kat:
        m3 = spark_ml.linear_regression_fit([go], po, lab_params(reg_param=0.0, elastic_net_param=0.0))
        assert rel_err(m3.coefficients[0], kat["ols"]["coefficient"]) < RTOL
        assert rel_err(m3.intercept, kat["ols"]["intercept"]) < RTOL
        assert m3.objective_history == [0.0]


@pytest.mark.parametrize("name", NAMES + ["synth_small"])
def test_fits_match_golden_and_independent_solvers(name, golden):
    doc = golden(name if name == "synth_small" else "dataset_" + name)
    if name == "synth_small":
        guest, price, extra = passes.synth_generate(doc["n"], d=doc["d"], p_low=doc["p_low"], p_high=doc["p_high"])
        assert guest[:16].tolist() == doc["head_guest"] and price[:16].tolist() == doc["head_price"]
        assert [str(v) for v in passes.rows_digest(guest, price)] == doc["digest"]
        keep = np.array([bool(passes.lib().orc_dq_keep(int(a), float(b))) for a, b in zip(guest, price)])
        feats = [guest[keep]] + [np.ascontiguousarray(extra[k][keep]) for k in range(doc["d"] - 1)]
        label = price[keep]
        assert int(keep.sum()) == doc["n_survivors"]
    else:
        feats = [np.asarray(doc["survivor_guest"], dtype=np.int32)]
        label = np.asarray(doc["survivor_price"], dtype=np.float64)
        cf = doc["closed_form_lasso"]
        lab = doc["fits"]["lab_wls_lasso"]
        assert rel_err(lab["coefficients"][0], cf["coefficient"]) < RTOL and rel_err(lab["intercept"], cf["intercept"]) < RTOL
    for fit_name, ref in doc["fits"].items():
        m = spark_ml.linear_regression_fit(feats, label, spark_ml.LRParams(**ref["params"]))
        assert m.solver_used == ref["solver_used"], fit_name
        assert rel_err(m.coefficients, ref["coefficients"]) < 1e-12, fit_name
        assert rel_err(m.intercept, ref["intercept"]) < 1e-12, fit_name
        assert rel_err(m.objective_history, ref["objective_history"]) < 1e-12, fit_name
        if "sklearn_coefficients" in ref:
            # independent converged optimum; OWL-QN / L-BFGS stop at tol = 1e-6, so compare at 1e-5 absolute-ish
            np.testing.assert_allclose(m.coefficients, ref["sklearn_coefficients"], rtol=2e-5, atol=2e-6, err_msg=fit_name)
            np.testing.assert_allclose(m.intercept, ref["sklearn_intercept"], rtol=2e-5, err_msg=fit_name)


def test_rules_edge_cases():
    nan, inf = float("nan"), float("inf")
    assert passes.check_minimum_price(19.999999) == -1 and passes.check_minimum_price(20.0) == 20.0
    assert np.isnan(passes.check_minimum_price(nan)) and passes.check_minimum_price(-0.0) == -1
    assert passes.check_price_range(90.0, 13) == 90.0 and passes.check_price_range(90.0000001, 13) == -1
    assert passes.check_price_range(500.0, 14) == 500.0 and np.isnan(passes.check_price_range(nan, 1))
    L = passes.lib()
    # Spark SQL ordering: NaN > 0 is true; +inf survives when guest >= 14; -inf, 0, -0 do not
    assert L.orc_dq_keep(1, nan) == 1 and L.orc_dq_keep(20, inf) == 1 and L.orc_dq_keep(5, inf) == 0
    assert L.orc_dq_keep(20, -inf) == 0 and L.orc_dq_keep(20, 0.0) == 0 and L.orc_dq_keep(13, 90.0) == 1
    # nulls: rule 2 maps them to -1 (CorrUdf:12-14), rule 1 throws (MinUdf:12)
    out = passes.udf_price_correlation([50.0, 50.0, 50.0], [20, 20, 20], price_valid=[1, 0, 1], guest_valid=[1, 1, 0])
    assert out.tolist() == [50.0, -1.0, -1.0]
    with pytest.raises(ValueError):
        passes.udf_minimum_price([50.0, 60.0], valid=[1, 0])


def test_empty_and_ragged_inputs():
    g0, p0 = np.zeros(0, dtype=np.int32), np.zeros(0, dtype=np.float64)
    go, po, n1 = passes.dq_filter(g0, p0)
    assert len(go) == 0 and n1 == 0
    assert passes.csv_parse(b"") [0].size == 0
    g, p = passes.csv_parse(b"1,23.24\r\n2,30\n\n3,5e1\r")
    assert g.tolist() == [1, 2, 3] and p.tolist() == [23.24, 30.0, 50.0]
    with pytest.raises(ValueError):
        passes.csv_parse(b"1;2\r")
    with pytest.raises(ValueError):
        spark_ml.wls_fit_from_moments(np.zeros(8), 1, spark_ml.LRParams())


def test_threaded_passes_agree_with_sequential_fold():
    guest, price, extra = passes.synth_generate(50_000, d=4)
    feats = [guest, extra[0], extra[1], extra[2]]
    a = passes.wls_moments(feats, price, dq=(guest, price), threads=1)
    b = passes.wls_moments(feats, price, dq=(guest, price), threads=4)
    assert a[0] == b[0] and rel_err(a, b) < 1e-12
    s1 = passes.summarizer(feats, price, dq=(guest, price), threads=1)
    s4 = passes.summarizer(feats, price, dq=(guest, price), threads=4)
    assert s1["count"] == s4["count"] and rel_err(s1["variance"], s4["variance"]) < 1e-12
    # summarizer vs numpy on the survivors
    keep = np.array([bool(passes.lib().orc_dq_keep(int(x), float(y))) for x, y in zip(guest, price)])
    assert rel_err(s1["mean"][-1], price[keep].mean()) < 1e-12
    assert rel_err(s1["variance"][1], extra[0][keep].var(ddof=1)) < 1e-11


@pytest.mark.parametrize("solver,reg,en", [("normal", 0.0, 0.0), ("normal", 0.3, 0.5), ("l-bfgs", 0.0, 0.0)])
def test_weighted_restatement_equals_replicated_rows(solver, reg, en):
    """setWeightCol: integer weights must act like repeated rows -- exactly for the normal-equation
    branch (population moments) and for the unregularised l-bfgs branch (the optimum does not depend on the scaling)."""
    rng = np.random.default_rng(1)
    n = 400
    g = rng.integers(1, 40, n).astype(np.int32)
    x2 = rng.normal(size=n)
    y = 20 + 5 * g + 3 * x2 + rng.normal(size=n)
    w = rng.integers(0, 4, n).astype(np.float64)
    idx = np.repeat(np.arange(n), w.astype(np.int64))
    p = spark_ml.LRParams(max_iter=100, reg_param=reg, elastic_net_param=en, tol=1e-10, fit_intercept=True,
                          standardization=True, solver=solver)
    mw = spark_ml.linear_regression_fit([g, x2], y, p, weight=w, dq=(g, y))
    mr = spark_ml.linear_regression_fit([g[idx], x2[idx]], y[idx], p, dq=(g[idx], y[idx]))
    np.testing.assert_allclose(mw.coefficients, mr.coefficients, rtol=2e-7)
    assert abs(mw.intercept - mr.intercept) < 2e-6
    s = spark_ml.summarizer_weighted([g, x2], y, w)
    on = w != 0
    assert s["count"] == on.sum()
    np.testing.assert_allclose(s["mean"][2], np.average(y[on], weights=w[on]), rtol=1e-13)
    with pytest.raises(ValueError):
        spark_ml.wls_moments_weighted([g], y, -w - 1.0)
