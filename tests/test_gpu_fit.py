"""GPU parity, ML section: aggregation passes and whole fits against the oracle, through the C ABI."""
import numpy as np
import pytest
import torch

from conftest import RTOL, rel_err, scale_err
from oracle import passes, spark_ml

pytestmark = pytest.mark.gpu

# Wide fits (d = 40 ... 256) are held to the same 1e-9 as everything else: the largest error measured on any fit of this
# file is 7e-12 (profiles/r2_parity_errors.jsonl).
TOL_WIDE = RTOL


def to_dev(*arrs):
    return [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in arrs]


def synth(n, d, **kw):
    guest, price, extra = passes.synth_generate(n, d=d, **kw)
    feats = [guest] + [np.ascontiguousarray(extra[k]) for k in range(d - 1)]
    return guest, price, feats


def estimator(pkg, p):
    lr = pkg.LinearRegression().setMaxIter(p.get("max_iter", 100)).setRegParam(p.get("reg_param", 0.0))
    lr.setElasticNetParam(p.get("elastic_net_param", 0.0)).setTol(p.get("tol", 1e-6))
    lr.setFitIntercept(p.get("fit_intercept", True)).setStandardization(p.get("standardization", True))
    lr.setSolver(p.get("solver", "auto"))
    return lr


def check_fit(parity_log, tag, model, ref_coef, ref_intercept, ref_hist, tol):
    """Whole-fit parity in the measure DESIGN.md states: objective history and intercept relative, coefficients relative
    to the largest coefficient (an L1 term leaves some at or next to zero, where element-wise relative error means
    nothing).  The measured errors go to the parity log."""
    e_hist = rel_err(model.summary().objectiveHistory(), ref_hist)
    e_coef = scale_err(model.coefficients(), ref_coef)
    e_int = rel_err(model.intercept(), ref_intercept)
    parity_log(tag, "objectiveHistory rel / coef scale-rel / intercept rel", [e_hist, e_coef, e_int])
    assert e_hist < tol and e_coef < tol and e_int < tol, (tag, e_hist, e_coef, e_int)


@pytest.mark.parametrize("d", [1, 2, 3, 4, 7, 8])
@pytest.mark.parametrize("n", [1, 5, 1023, 4096, 200_003])
def test_wls_moments_vs_oracle(pkg, ctx, d, n):
    from sparkdq4ml_b200 import ml
    guest, price, feats = synth(n, d, p_low=0.05, p_high=0.05)
    ref = passes.wls_moments(feats, price, dq=(guest, price), threads=4)
    dfeats = to_dev(*feats)
    dprice = to_dev(price)[0]
    got = ml.wls_moments(ctx, dfeats, dprice, dq=(dfeats[0], dprice))  # lab layout: guest is feature 0, price is label
    assert got[0] == ref[0] and got[1] == ref[1]  # survivor count is exact
    assert rel_err(got, ref) < 1e-11
    got2 = ml.wls_moments(ctx, dfeats, dprice, dq=(dfeats[0].clone(), dprice.clone()))  # DQ columns passed separately
    # same rows, same rules; at d = 1 the lab layout takes the software-pipelined kernel, whose threads own other rows
    # than the grid-strided kernel's, so the sums may differ in the last bit
    assert got2[0] == got[0] and rel_err(got2, got) < 1e-13
    if d > 1:
        assert got2.tobytes() == got.tobytes()
    no_dq = ml.wls_moments(ctx, dfeats, dprice)
    assert rel_err(no_dq, passes.wls_moments(feats, price, threads=4)) < 1e-11
    host = ml.wls_moments(ctx, feats, price, dq=(guest, price))  # host buffers through the same entry point
    assert host.tobytes() == got.tobytes()


@pytest.mark.parametrize("n", [6143, 6144, 6145, 12_288, 148 * 6144 - 1, 148 * 6144, 148 * 6144 + 1, 149 * 6144 + 777])
def test_pipelined_single_column_kernel_tile_boundaries(pkg, ctx, n):
    """The lab's shape (one int32 feature + label, DQ in the lab layout or off) takes the software-pipelined kernel:
    6144-row tiles, one CTA per SM (148 on a B200), the ragged end row by row.  Moments, evaluation sums and the per-row
    residuals on both sides of every boundary, against the oracle."""
    from sparkdq4ml_b200 import ml
    guest, price, feats = synth(n, 1, p_low=0.03, p_high=0.03)
    dguest, dprice = to_dev(guest, price)
    for dq_host, dq_dev in (((guest, price), (dguest, dprice)), (None, None)):
        ref = passes.wls_moments(feats, price, dq=dq_host, threads=4)
        got = ml.wls_moments(ctx, [dguest], dprice, dq=dq_dev)
        assert got[0] == ref[0] and rel_err(got, ref) < 1e-11
        assert ml.wls_moments(ctx, [dguest], dprice, dq=dq_dev).tobytes() == got.tobytes()  # fixed-order reduction
    cols = {"guest": dguest, "label": dprice}
    df = pkg.VectorAssembler().setInputCols(["guest"]).transform(pkg.Frame(ctx, cols).with_fused_dq("guest", "label"))
    model = pkg.LinearRegression().setMaxIter(40).setRegParam(1.0).setElasticNetParam(1.0).fit(df)
    keep = np.array([bool(passes.lib().orc_dq_keep(int(a), float(b))) for a, b in zip(guest, price)])
    res_ref = price - (guest * model.coefficients()[0] + model.intercept())
    ev = model.evaluate(df, want_residuals=True)
    res = ev.residuals().cpu().numpy()
    np.testing.assert_allclose(res[keep], res_ref[keep], rtol=0, atol=1e-9 * float(np.abs(price).max()))
    assert np.all(res[~keep] == 0.0) and ev.numInstances() == int(keep.sum())
    assert rel_err(ev.rootMeanSquaredError(), float(np.sqrt(np.mean(res_ref[keep] ** 2)))) < 1e-10


def test_wls_moments_empty_and_all_dropped(pkg, ctx):
    from sparkdq4ml_b200 import ml
    g = torch.zeros(0, dtype=torch.int32, device="cuda")
    p = torch.zeros(0, dtype=torch.float64, device="cuda")
    assert ml.wls_moments(ctx, [g], p).tolist() == [0.0] * 8
    g = torch.full((1000,), 3, dtype=torch.int32, device="cuda")
    p = torch.full((1000,), 5.0, dtype=torch.float64, device="cuda")
    assert ml.wls_moments(ctx, [g], p, dq=(g, p)).tolist() == [0.0] * 8
    df = pkg.Frame(ctx, {"guest": g, "price": p, "label": p}).with_fused_dq()
    df = pkg.VectorAssembler().setInputCols(["guest"]).transform(df)
    with pytest.raises(pkg.capi.Dq4mlError) as ei:
        pkg.LinearRegression().fit(df)
    assert ei.value.status == pkg.capi.ERR_EMPTY


def test_nan_price_survives_dq_and_poisons_moments_like_spark(pkg, ctx):
    from sparkdq4ml_b200 import ml
    g = np.array([20, 21, 22, 23, 5], dtype=np.int32)
    p = np.array([100.0, float("nan"), 110.0, 10.0, 120.0])
    ref = passes.wls_moments([g], p, dq=(g, p))
    got = ml.wls_moments(ctx, to_dev(g), to_dev(p)[0], dq=(to_dev(g)[0], to_dev(p)[0]))
    assert got[0] == ref[0] == 3.0
    assert np.isnan(got[3]) and np.isnan(ref[3]) and got[5] == ref[5]


@pytest.mark.parametrize("name", ["abstract", "small", "full"])
@pytest.mark.parametrize("fused", [False, True])
def test_lab_app_on_reference_fixtures(pkg, ctx, golden, lab_csv, name, fused):
    """Config 1: DataQuality4MachineLearningApp on the reference's CSVs."""
    kat, doc = golden("survey_kats")[name], golden("dataset_" + name)
    out = pkg.DataQuality4MachineLearningApp(ctx).start(lab_csv(name), fused=fused)
    ref = doc["fits"]["lab_wls_lasso"]
    w = kat["wls_lasso"]
    assert out["n_loaded"] == kat["n"] and out["n_fit_rows"] == kat["n_after_rule2"]
    if not fused:
        assert (out["n_after_rule1"], out["n_after_rule2"]) == (kat["n_after_rule1"], kat["n_after_rule2"])
        assert out["survivor_guest"].cpu().numpy().tolist() == doc["survivor_guest"]
        assert out["survivor_price"].cpu().numpy().tolist() == doc["survivor_price"]
    assert out["solver_used"] == "wls-owlqn" and out["numIterations"] == w["n_states"] == ref["total_iterations"]
    assert rel_err(out["coefficients"], ref["coefficients"]) < RTOL and rel_err(out["coefficients"][0], w["coefficient"]) < RTOL
    assert rel_err(out["intercept"], ref["intercept"]) < RTOL and rel_err(out["intercept"], w["intercept"]) < RTOL
    assert rel_err(out["objectiveHistory"], ref["objective_history"]) < RTOL
    assert rel_err(out["prediction"], w["predict_40"]) < RTOL
    assert rel_err(out["RMSE"], w["rmse"]) < RTOL and rel_err(out["r2"], w["r2"]) < RTOL
    assert out["regParam"] == 1.0 and out["tol"] == 1e-6
    # per-row predictions and residuals against the oracle
    go = np.asarray(doc["survivor_guest"], dtype=np.int32)
    po = np.asarray(doc["survivor_price"], dtype=np.float64)
    pred_ref, _ = passes.predict_metrics([go], po, np.asarray(ref["coefficients"]), ref["intercept"])
    pred = out["prediction_column"].cpu().numpy()
    res = out["residuals"].cpu().numpy()
    if fused:  # rows the rules drop stay in place: compare on the survivors
        g_all = np.asarray(doc["guest"], dtype=np.int32)
        p_all = np.asarray(doc["price"], dtype=np.float64)
        keep = np.array([bool(passes.lib().orc_dq_keep(int(a), float(b))) for a, b in zip(g_all, p_all)])
        assert (res[~keep] == 0).all()
        pred, res = pred[keep], res[keep]
    assert rel_err(pred, pred_ref) < RTOL
    np.testing.assert_allclose(res, po - pred_ref, rtol=0, atol=1e-9 * np.abs(po).max())


@pytest.mark.parametrize("fit_name", ["lab_wls_lasso", "ols_cholesky", "ridge_cholesky", "enet_wls", "nostd_enet_wls",
                                      "lab_lbfgs_lasso", "enet_lbfgs", "ols_lbfgs", "nointercept_lbfgs"])
def test_fit_synth_small_all_solvers(pkg, ctx, golden, parity_log, fit_name):
    """d = 4 with fused DQ: every solver branch against the golden fits (oracle) and scikit-learn."""
    doc = golden("synth_small")
    ref = doc["fits"][fit_name]
    g, p, x = ctx.synth_generate(doc["n"], d=doc["d"], p_low=doc["p_low"], p_high=doc["p_high"])
    cols = {"guest": g, "price": p, "label": p}
    names = ["guest"]
    for k in range(doc["d"] - 1):
        cols[f"x{k + 1}"] = x[k]
        names.append(f"x{k + 1}")
    df = pkg.Frame(ctx, cols).with_fused_dq("guest", "price")
    df = pkg.VectorAssembler().setInputCols(names).setOutputCol("features").transform(df)
    model = estimator(pkg, ref["params"]).fit(df)
    assert model.num_rows == doc["n_survivors"]
    assert model.solver_used == ref["solver_used"]
    s = model.summary()
    assert s.totalIterations() == ref["total_iterations"]
    if abs(ref["intercept"]) > 0:
        check_fit(parity_log, "synth_small/" + fit_name, model, ref["coefficients"], ref["intercept"], ref["objective_history"], RTOL)
    else:  # fitIntercept = false
        assert model.intercept() == 0.0
        assert rel_err(s.objectiveHistory(), ref["objective_history"]) < RTOL
        assert scale_err(model.coefficients(), ref["coefficients"]) < RTOL
    assert rel_err(s.rootMeanSquaredError(), ref["rmse"]) < RTOL and rel_err(s.r2(), ref["r2"]) < RTOL
    if "sklearn_coefficients" in ref:
        np.testing.assert_allclose(model.coefficients(), ref["sklearn_coefficients"], rtol=2e-5, atol=2e-6)


# 9..12: lane-group kernels; 13..16: generic bucket; 17..24: lane pairs; 25..32: groups of four lanes
@pytest.mark.parametrize("d", [1, 3, 8, 9, 10, 12, 13, 16, 17, 20, 24, 25, 28, 32])
def test_summarizer_and_lsq_passes_vs_oracle(pkg, ctx, d):
    from sparkdq4ml_b200 import ml
    n = 150_001
    guest, price, feats = synth(n, d, p_low=0.04, p_high=0.04)
    dfeats, dprice = to_dev(*feats), to_dev(price)[0]
    ref = passes.summarizer(feats, price, dq=(guest, price), threads=4)
    got = ml.summarize(ctx, dfeats, dprice, dq=(dfeats[0], dprice))
    assert got["count"] == ref["count"]
    for key in ("mean", "variance", "normL1", "normL2"):
        assert rel_err(got[key], ref[key]) < 1e-10, key
    for key in ("min", "max", "nnz"):
        assert got[key].tolist() == ref[key].tolist(), key
    rng = np.random.default_rng(d)
    coef = rng.normal(size=d)
    fstd, fmean = np.sqrt(ref["variance"][:d]), ref["mean"][:d]
    ystd, ymean = np.sqrt(ref["variance"][d]), ref["mean"][d]
    for fit_intercept in (True, False):
        rl, rw, rg = passes.lsq_loss_grad(feats, price, coef, fstd, fmean, ystd, ymean, fit_intercept, dq=(guest, price), threads=4)
        gl, gw, gg = ml.lsq_loss_grad(ctx, dfeats, dprice, coef, fstd, fmean, ystd, ymean, fit_intercept, dq=(dfeats[0], dprice))
        assert gw == rw
        assert rel_err(gl, rl) < 1e-10
        np.testing.assert_allclose(gg, rg, rtol=1e-9, atol=1e-9 * np.abs(rg).max())


@pytest.mark.parametrize("reg_param", [1.0, 0.3])
def test_config3_shape_d16_owlqn_fused(pkg, ctx, parity_log, reg_param):
    """BASELINE config 3 at a test-sized n: 16 features through VectorAssembler, l-bfgs solver, elasticNet 0.5;
    regParam 1 is what the benchmark runs (tests/test_gpu_multirank.py repeats it at 10^7 rows), 0.3 keeps more
    coefficients alive."""
    n, d = 400_000, 16
    guest, price, feats = synth(n, d)
    params = dict(max_iter=40, reg_param=reg_param, elastic_net_param=0.5, solver="l-bfgs")
    ref = spark_ml.linear_regression_fit(feats, price, spark_ml.LRParams(**params), dq=(guest, price), threads=8)
    g, p, x = ctx.synth_generate(n, d=d)
    cols = {"guest": g, "label": p}
    names = ["guest"]
    for k in range(d - 1):
        cols[f"x{k + 1}"] = x[k]
        names.append(f"x{k + 1}")
    df = pkg.Frame(ctx, cols).with_fused_dq("guest", "label")
    df = pkg.VectorAssembler().setInputCols(names).transform(df)
    model = estimator(pkg, params).fit(df)
    assert model.solver_used == ref.solver_used == "owlqn"
    s = model.summary()
    assert s.totalIterations() == ref.total_iterations and model.data_passes == ref.data_passes
    check_fit(parity_log, f"config3_shape_reg{reg_param}", model, ref.coefficients, ref.intercept, ref.objective_history, RTOL)


@pytest.mark.parametrize("d,n", [(8, 20_001), (9, 30_001), (10, 4_095), (12, 50_000), (14, 20_003), (16, 40_003), (20, 30_000), (24, 12_289),
                                 (28, 20_000), (32, 25_001)])
def test_lbfgs_fit_and_evaluation_by_column_bucket(pkg, ctx, parity_log, d, n):
    """One l-bfgs fit and one evaluation per kernel family of the streaming passes (full buckets of 8 / 16 / 32 columns:
    generic kernel with the compile-time load phase; narrower tables: compile-time column types for the dot-product ops,
    lane-group / lane-pair kernels for the summarizer and for 17..20 columns): the fit against the oracle, the
    evaluation's per-row residuals and metrics against numpy on the surviving rows."""
    guest, price, feats = synth(n, d, p_low=0.03, p_high=0.03)
    params = dict(max_iter=30, reg_param=0.2, elastic_net_param=0.5, solver="l-bfgs")
    ref = spark_ml.linear_regression_fit(feats, price, spark_ml.LRParams(**params), dq=(guest, price), threads=4)
    dfeats, dprice = to_dev(*feats), to_dev(price)[0]
    cols = {"guest": dfeats[0], "label": dprice}
    names = ["guest"]
    for k in range(1, d):
        cols[f"x{k}"] = dfeats[k]
        names.append(f"x{k}")
    df = pkg.VectorAssembler().setInputCols(names).transform(pkg.Frame(ctx, cols).with_fused_dq("guest", "label"))
    model = estimator(pkg, params).fit(df)
    assert model.solver_used == ref.solver_used == "owlqn"
    assert model.summary().totalIterations() == ref.total_iterations and model.data_passes == ref.data_passes
    check_fit(parity_log, f"bucket_d{d}", model, ref.coefficients, ref.intercept, ref.objective_history, RTOL)
    keep = np.array([bool(passes.lib().orc_dq_keep(int(a), float(b))) for a, b in zip(guest, price)])
    X = np.stack([np.asarray(f, dtype=np.float64) for f in feats], axis=1)
    res_ref = price - (X @ model.coefficients() + model.intercept())
    ev = model.evaluate(df, want_residuals=True)
    res = ev.residuals().cpu().numpy()
    np.testing.assert_allclose(res[keep], res_ref[keep], rtol=0, atol=1e-9 * max(1.0, float(np.abs(price).max())))
    assert ev.numInstances() == int(keep.sum())
    assert rel_err(ev.rootMeanSquaredError(), float(np.sqrt(np.mean(res_ref[keep] ** 2)))) < 1e-9
    assert rel_err(ev.meanAbsoluteError(), float(np.mean(np.abs(res_ref[keep])))) < 1e-9
    lo, hi = ev.devianceResiduals()
    assert rel_err(lo, float(res_ref[keep].min())) < 1e-9 and rel_err(hi, float(res_ref[keep].max())) < 1e-9


def test_host_buffers_equal_device_buffers(pkg, ctx):
    n = 50_000
    guest, price, feats = synth(n, 3)
    def fit(cols_g, cols_p, cols_x):
        df = pkg.Frame(ctx, {"guest": cols_g, "price": cols_p, "label": cols_p, "x1": cols_x[0], "x2": cols_x[1]}).with_fused_dq()
        df = pkg.VectorAssembler().setInputCols(["guest", "x1", "x2"]).transform(df)
        m = pkg.LinearRegression().setMaxIter(40).setRegParam(0.2).setElasticNetParam(1.0).fit(df)
        return m.coefficients(), m.intercept(), m.summary().rootMeanSquaredError()
    host = fit(guest, price, feats[1:])
    dev = fit(*to_dev(guest, price), to_dev(*feats[1:]))
    assert host[0].tobytes() == dev[0].tobytes() and host[1] == dev[1] and host[2] == dev[2]


def test_repeatability_bitwise(pkg, ctx):
    from sparkdq4ml_b200 import ml
    g, p, _ = ctx.synth_generate(3_000_001, d=1)
    a = ml.wls_moments(ctx, [g], p, dq=(g, p))
    b = ml.wls_moments(ctx, [g], p, dq=(g, p))
    assert a.tobytes() == b.tobytes()


def test_fit_at_config2_size_properties(pkg, ctx):
    """BASELINE config 2 size (1e8 rows, d = 1, lasso): checked through size-independent properties + the oracle's fit."""
    from sparkdq4ml_b200 import ml
    n = 100_000_000
    g, p, _ = ctx.synth_generate(n, d=1)
    mom = ml.wls_moments(ctx, [g], p, dq=(g, p))
    # linearity: moments of the whole = sum of the moments of two halves (shard boundary on a 1024-row multiple)
    h = (n // 2) // 1024 * 1024
    a = ml.wls_moments(ctx, [g[:h]], p[:h], dq=(g[:h], p[:h]))
    b = ml.wls_moments(ctx, [g[h:]], p[h:], dq=(g[h:], p[h:]))
    assert mom[0] == a[0] + b[0]
    assert rel_err(mom, a + b) < 1e-12
    # filter-then-aggregate == fused
    go, po, _ = pkg.dq_filter(ctx, g, p)
    assert go.shape[0] == mom[0]
    assert rel_err(ml.wls_moments(ctx, [go], po), mom) < 1e-12
    og, op, _ = passes.synth_generate(n, d=1)
    ref_mom = passes.wls_moments([og], op, dq=(og, op), threads=passes.max_threads())
    assert ref_mom[0] == mom[0] and rel_err(mom, ref_mom) < 1e-11
    lr = pkg.LinearRegression().setMaxIter(40).setRegParam(1).setElasticNetParam(1)
    df = pkg.VectorAssembler().setInputCols(["guest"]).transform(pkg.Frame(ctx, {"guest": g, "price": p, "label": p}).with_fused_dq())
    model = lr.fit(df)
    ref = spark_ml.wls_fit_from_moments(ref_mom, 1, spark_ml.LRParams(max_iter=40, reg_param=1.0, elastic_net_param=1.0))
    assert rel_err(model.coefficients(), ref.coefficients) < RTOL and rel_err(model.intercept(), ref.intercept) < RTOL
    assert rel_err(model.summary().objectiveHistory(), ref.objective_history) < RTOL


# ---------------------------------------------------------------------------------------------- K5: wide tables (Gram)

def wide_table(n, d, seed=0, int_cols=(0,)):
    """guest (int32) + d-1 synthetic f64 columns; label = price."""
    guest, price, extra = passes.synth_generate(n, d=d, p_low=0.05, p_high=0.05)
    feats = [guest] + [np.ascontiguousarray(extra[k]) for k in range(d - 1)]
    return guest, price, feats


@pytest.mark.parametrize("d,n", [(9, 1), (9, 31), (16, 32), (16, 4097), (33, 1000), (64, 50_001), (65, 20_000),
                                 (130, 10_007), (256, 20_011)])
def test_gram_wls_moments_vs_oracle(pkg, ctx, d, n):
    from sparkdq4ml_b200 import ml
    guest, price, feats = wide_table(n, d)
    ref = passes.wls_moments(feats, price, dq=(guest, price), threads=8)
    dfeats, dprice = to_dev(*feats), to_dev(price)[0]
    got = ml.wls_moments(ctx, dfeats, dprice, dq=(dfeats[0], dprice))
    assert got.shape == ref.shape
    assert got[0] == ref[0] == got[1] == got[2]  # survivor count exact
    scale = np.maximum(np.abs(ref), 1e-300)
    assert float(np.max(np.abs(got - ref) / scale)) < 1e-11
    again = ml.wls_moments(ctx, dfeats, dprice, dq=(dfeats[0], dprice))
    assert again.tobytes() == got.tobytes()  # fixed-order reduction: bit-reproducible
    nodq = ml.wls_moments(ctx, dfeats, dprice)
    ref2 = passes.wls_moments(feats, price, threads=8)
    assert float(np.max(np.abs(nodq - ref2) / np.maximum(np.abs(ref2), 1e-300))) < 1e-11
    host = ml.wls_moments(ctx, feats, price, dq=(guest, price))  # host columns through the same entry point
    assert host.tobytes() == got.tobytes()


def test_gram_nan_and_inf_rows(pkg, ctx):
    """Rows the rules drop may hold anything (-inf price is dropped by rule 1) and must not poison the sums; a NaN price
    survives DQ (Spark ordering) and must poison exactly the sums Spark's would."""
    from sparkdq4ml_b200 import ml
    d, n = 12, 300
    guest, price, feats = wide_table(n, d)
    price = price.copy()
    price[7] = -np.inf
    feats[3] = feats[3].copy()
    feats[3][7] = np.nan  # garbage in a dropped row
    ref = passes.wls_moments(feats, price, dq=(guest, price))
    got = ml.wls_moments(ctx, to_dev(*feats), to_dev(price)[0], dq=(to_dev(guest)[0], to_dev(price)[0]))
    assert np.isfinite(ref).all() and np.isfinite(got).all()
    assert float(np.max(np.abs(got - ref) / np.maximum(np.abs(ref), 1e-300))) < 1e-11
    price[11] = np.nan
    ref = passes.wls_moments(feats, price, dq=(guest, price))
    got = ml.wls_moments(ctx, to_dev(*feats), to_dev(price)[0], dq=(to_dev(guest)[0], to_dev(price)[0]))
    assert (np.isnan(ref) == np.isnan(got)).all() and got[0] == ref[0]


def test_gram_dq_predicate_on_edge_prices(pkg, ctx):
    """The Gram kernel's producer warps evaluate the two rules on the bit pattern of the price (no fp64 compare): every
    edge of the IEEE ordering must come out as the oracle's fp64 compares do -- signed zeros, the thresholds
    themselves and their neighbours, denormals, infinities, NaN with either sign."""
    from sparkdq4ml_b200 import ml
    d, n = 12, 4096 + 17
    guest, price, feats = wide_table(n, d)
    price = price.copy()
    guest = guest.copy()
    edges = [0.0, -0.0, 5e-324, -5e-324, 20.0, np.nextafter(20.0, 0), np.nextafter(20.0, 100), 90.0, np.nextafter(90.0, 0),
             np.nextafter(90.0, 100), np.inf, -np.inf, np.nan, -np.nan, 1e308, -1e308, 2.2250738585072014e-308, 19.999999999, 90.000001]
    k = 0
    for g in (10, 13, 14, 15, 40):  # min_guest is 14: both sides of the correlation rule
        for e in edges:
            price[k * 7 + 3] = e
            guest[k * 7 + 3] = g
            k += 1
    feats[0] = guest
    from sparkdq4ml_b200 import _capi as capi
    dfeats, dprice, dguest = to_dev(*feats), to_dev(price)[0], to_dev(guest)[0]
    ref = passes.wls_moments(feats, price, dq=(guest, price))
    got = ml.wls_moments(ctx, dfeats, dprice, dq=(dguest, dprice))
    assert got[0] == ref[0]  # survivor count: the predicate agrees row for row on the planted edges
    assert (np.isnan(ref) == np.isnan(got)).all()
    # other thresholds (the oracle has the lab's constants only): the same truth table in numpy, IEEE compares
    for min_price, max_price, min_guest in ((25.5, 60.25, 12), (-0.0, np.inf, 0), (0.0, 90.0, 14), (np.nan, np.nan, 14), (1e-310, 1e300, 41)):
        with np.errstate(invalid="ignore"):
            keep = ~(price < min_price) & ~(price <= 0.0) & ~((guest < min_guest) & (price > max_price))
        ref = passes.wls_moments([np.ascontiguousarray(f[keep]) for f in feats], np.ascontiguousarray(price[keep]))
        got = ml.wls_moments(ctx, dfeats, dprice, dq=(dguest, dprice), dq_params=capi.DqParams(min_price, min_guest, max_price))
        assert got[0] == ref[0] == float(keep.sum())
        assert (np.isnan(ref) == np.isnan(got)).all()
        ok = ~np.isnan(ref)
        assert float(np.max(np.abs(got[ok] - ref[ok]) / np.maximum(np.abs(ref[ok]), 1e-300))) < 1e-11


@pytest.mark.parametrize("params", [dict(max_iter=40, reg_param=0.0, elastic_net_param=0.0, solver="normal"),
                                    dict(max_iter=40, reg_param=0.1, elastic_net_param=0.0),
                                    dict(max_iter=40, reg_param=0.05, elastic_net_param=0.5)])
def test_config4_shape_d256_normal_equations(pkg, ctx, parity_log, params):
    """BASELINE config 4 at a test-sized n: 256 features, normal-equation solver (Cholesky, or OWL-QN over the Gram when
    an L1 term is present), DQ fused."""
    n, d = 60_000, 256
    guest, price, feats = wide_table(n, d)
    ref = spark_ml.linear_regression_fit(feats, price, spark_ml.LRParams(**params), dq=(guest, price), threads=8)
    g, p, x = ctx.synth_generate(n, d=d, p_low=0.05, p_high=0.05)
    cols = {"guest": g, "label": p}
    names = ["guest"]
    for k in range(d - 1):
        cols[f"x{k + 1}"] = x[k]
        names.append(f"x{k + 1}")
    df = pkg.Frame(ctx, cols).with_fused_dq("guest", "label")
    df = pkg.VectorAssembler().setInputCols(names).transform(df)
    model = estimator(pkg, params).fit(df)
    assert model.solver_used == ref.solver_used and model.num_rows == int(passes.wls_moments([guest], price, dq=(guest, price))[0])
    assert model.summary().totalIterations() == ref.total_iterations
    check_fit(parity_log, f"config4_shape_{params['solver'] if 'solver' in params else 'auto'}_reg{params['reg_param']}_en{params['elastic_net_param']}",
              model, ref.coefficients, ref.intercept, ref.objective_history, TOL_WIDE)


# ---------------------------------------------------------------------------------------------- wide tables beyond the Gram

@pytest.mark.parametrize("d,n", [(33, 5_001), (48, 70_000), (256, 30_003)])
def test_wide_transform_evaluate_and_lsq_vs_oracle(pkg, ctx, d, n):
    """Tables wider than the 32 register-resident columns: column-walking predict/evaluate pass and the two-pass
    loss/gradient, fused DQ, against the oracle."""
    from sparkdq4ml_b200 import ml
    guest, price, feats = wide_table(n, d)
    rng = np.random.default_rng(d)
    coef = rng.normal(size=d) * 0.1
    intercept = 3.25
    pred_ref, met_ref = passes.predict_metrics(feats, price, coef, intercept, dq=(guest, price), threads=8)
    dfeats, dprice = to_dev(*feats), to_dev(price)[0]
    cols = {"guest": dfeats[0], "label": dprice}
    names = ["guest"]
    for k in range(1, d):
        cols[f"x{k}"] = dfeats[k]
        names.append(f"x{k}")
    df = pkg.VectorAssembler().setInputCols(names).transform(pkg.Frame(ctx, cols).with_fused_dq("guest", "label"))
    model = pkg.LinearRegressionModel.from_coefficients(coef, intercept)
    pred = model.transform(df).col("prediction").cpu().numpy()
    keep = np.array([bool(passes.lib().orc_dq_keep(int(a), float(b))) for a, b in zip(guest, price)])
    plain, _ = passes.predict_metrics(feats, price, coef, intercept, threads=8)  # transform ignores DQ
    np.testing.assert_allclose(pred, plain, rtol=1e-12, atol=1e-10)
    ev = model.evaluate(df, want_residuals=True)
    ref = spark_ml.regression_metrics(met_ref)
    assert ev.numInstances() == int(met_ref[0])
    assert rel_err(ev.rootMeanSquaredError(), ref["rmse"]) < 1e-10 and rel_err(ev.r2(), ref["r2"]) < 1e-9
    assert rel_err(ev.meanAbsoluteError(), ref["mae"]) < 1e-10
    res = ev.residuals().cpu().numpy()
    assert (res[~keep] == 0).all()
    np.testing.assert_allclose(res[keep], (price - pred_ref)[keep], rtol=0, atol=1e-9 * np.abs(price).max())
    # loss / gradient
    summ = passes.summarizer(feats, price, dq=(guest, price), threads=8)
    fstd, fmean = np.sqrt(summ["variance"][:d]), summ["mean"][:d]
    ystd, ymean = np.sqrt(summ["variance"][d]), summ["mean"][d]
    c = rng.normal(size=d)
    rl, rw, rg = passes.lsq_loss_grad(feats, price, c, fstd, fmean, ystd, ymean, True, dq=(guest, price), threads=8)
    gl, gw, gg = ml.lsq_loss_grad(ctx, dfeats, dprice, c, fstd, fmean, ystd, ymean, True, dq=(dfeats[0], dprice))
    assert gw == rw and rel_err(gl, rl) < 1e-10
    np.testing.assert_allclose(gg, rg, rtol=1e-9, atol=1e-9 * np.abs(rg).max())


def test_wide_lbfgs_fit_d40(pkg, ctx, parity_log):
    """solver = l-bfgs on a 40-column table: grouped summarizer + two-pass loss/gradient under OWL-QN."""
    n, d = 50_000, 40
    guest, price, feats = wide_table(n, d)
    params = dict(max_iter=30, reg_param=0.2, elastic_net_param=0.5, solver="l-bfgs")
    ref = spark_ml.linear_regression_fit(feats, price, spark_ml.LRParams(**params), dq=(guest, price), threads=8)
    dfeats, dprice = to_dev(*feats), to_dev(price)[0]
    cols = {"guest": dfeats[0], "label": dprice}
    names = ["guest"]
    for k in range(1, d):
        cols[f"x{k}"] = dfeats[k]
        names.append(f"x{k}")
    df = pkg.VectorAssembler().setInputCols(names).transform(pkg.Frame(ctx, cols).with_fused_dq("guest", "label"))
    model = estimator(pkg, params).fit(df)
    assert model.solver_used == ref.solver_used == "owlqn"
    assert model.summary().totalIterations() == ref.total_iterations
    check_fit(parity_log, "wide_lbfgs_d40", model, ref.coefficients, ref.intercept, ref.objective_history, TOL_WIDE)


def test_wide_lbfgs_fit_d70_second_group_takes_the_lane_pair_kernel(pkg, ctx, parity_log):
    """70 columns under solver l-bfgs: the summarizer runs in groups of 32, 32 and 6 columns, and the second group -- more
    than 16 columns, no label -- goes through the lane-pair kernel without a label column."""
    n, d = 30_001, 70
    guest, price, feats = wide_table(n, d)
    params = dict(max_iter=25, reg_param=0.1, elastic_net_param=0.3, solver="l-bfgs")
    ref = spark_ml.linear_regression_fit(feats, price, spark_ml.LRParams(**params), dq=(guest, price), threads=8)
    dfeats, dprice = to_dev(*feats), to_dev(price)[0]
    cols = {"guest": dfeats[0], "label": dprice}
    names = ["guest"]
    for k in range(1, d):
        cols[f"x{k}"] = dfeats[k]
        names.append(f"x{k}")
    df = pkg.VectorAssembler().setInputCols(names).transform(pkg.Frame(ctx, cols).with_fused_dq("guest", "label"))
    model = estimator(pkg, params).fit(df)
    assert model.solver_used == ref.solver_used == "owlqn"
    assert model.summary().totalIterations() == ref.total_iterations
    check_fit(parity_log, "wide_lbfgs_d70", model, ref.coefficients, ref.intercept, ref.objective_history, TOL_WIDE)


@pytest.mark.parametrize("d", [3, 40])
def test_training_summary_surface_vs_numpy(pkg, ctx, d):
    """The rest of LinearRegressionTrainingSummary (explainedVariance, r2adj, degreesOfFreedom, devianceResiduals,
    standard errors, t- and p-values) on an OLS fit with fused DQ, against numpy/scipy on the surviving rows."""
    from scipy import stats
    n = 80_000
    guest, price, feats = wide_table(n, d)
    keep = np.array([bool(passes.lib().orc_dq_keep(int(a), float(b))) for a, b in zip(guest, price)])
    X = np.stack([np.asarray(f, dtype=np.float64)[keep] for f in feats], axis=1)
    y = price[keep]
    dfeats, dprice = to_dev(*feats), to_dev(price)[0]
    cols = {"guest": dfeats[0], "label": dprice}
    names = ["guest"]
    for k in range(1, d):
        cols[f"x{k}"] = dfeats[k]
        names.append(f"x{k}")
    df = pkg.VectorAssembler().setInputCols(names).transform(pkg.Frame(ctx, cols).with_fused_dq("guest", "label"))
    model = pkg.LinearRegression().setSolver("normal").fit(df)
    s = model.summary()
    m = len(y)
    Xa = np.hstack([X, np.ones((m, 1))])
    beta = np.linalg.lstsq(Xa, y, rcond=None)[0]
    pred = Xa @ beta
    res = y - pred
    np.testing.assert_allclose(np.append(model.coefficients(), model.intercept()), beta, rtol=1e-7, atol=1e-9)
    assert s.numInstances() == m and s.degreesOfFreedom() == m - d - 1
    r2 = 1 - (res @ res) / ((y - y.mean()) @ (y - y.mean()))
    assert rel_err(s.r2(), r2) < 1e-9
    assert rel_err(s.r2adj(), 1 - (1 - r2) * (m - 1) / (m - d - 1)) < 1e-9
    assert rel_err(s.explainedVariance(), ((pred - y.mean()) ** 2).mean()) < 1e-8
    assert rel_err(s.meanAbsoluteError(), np.abs(res).mean()) < 1e-8
    np.testing.assert_allclose(s.devianceResiduals(), [res.min(), res.max()], rtol=1e-6, atol=1e-7)
    dof = m - d - 1
    se_ref = np.sqrt(np.diag(np.linalg.inv(Xa.T @ Xa)) * (res @ res) / dof)
    np.testing.assert_allclose(s.coefficientStandardErrors(), se_ref, rtol=1e-6)
    np.testing.assert_allclose(s.tValues(), beta / se_ref, rtol=1e-6)
    np.testing.assert_allclose(s.pValues(), 2 * stats.t.sf(np.abs(beta / se_ref), dof), rtol=1e-5, atol=1e-300)
