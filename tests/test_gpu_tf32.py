"""GPU parity for the opt-in tcgen05 tier of the normal-equation pass (dq4ml_ctx_set_gram_tier, DQ4ML_GRAM_TF32X3).

Stated tolerance of the tier: 1e-5 relative on coefficients / intercept (BASELINE.json north_star: "1e-5 where fp32
accumulation is used and justified").  The moments are compared with the oracle's at the scale that matters for the
regression (sigma_i sigma_j for second moments), the fits with the oracle's fits."""
import numpy as np
import pytest
import torch

from oracle import passes, spark_ml

pytestmark = pytest.mark.gpu

COV_TOL = 5e-6   # |G - G_ref| / (n sigma_i sigma_j)
FIT_TOL = 1e-5   # the tier's stated tolerance: intercept relative; coefficients relative to the largest coefficient
# (measured on these cases: <= 2.2e-6 scale-relative, <= 4.2e-6 on the intercept, profiles/r2_parity_errors.jsonl)
COEF_FLOOR = 0.05      # per-coefficient relative error is checked for coefficients of at least 5 % of the largest one ...
PER_COEF_TOL = 1e-4    # ... where the tf32 Gram's absolute error is up to 20x larger in relative terms (measured 1.8e-5
                       # unweighted, 4.4e-5 with instance weights)


@pytest.fixture()
def tf32(ctx):
    ctx.set_gram_tier("tf32x3")
    yield ctx
    ctx.set_gram_tier("fp64")


def to_dev(*arrs):
    return [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in arrs]


def table(n, d):
    guest, price, extra = passes.synth_generate(n, d=d, p_low=0.05, p_high=0.05)
    feats = [guest] + [np.ascontiguousarray(extra[k]) for k in range(d - 1)]
    return guest, price, feats


def unpack(m, d):
    G = np.zeros((d, d))
    q = 0
    for j in range(d):
        for i in range(j + 1):
            G[i, j] = G[j, i] = m[5 + 2 * d + q]
            q += 1
    return G


def check_moments(got, ref, d):
    assert got[0] == ref[0]                                    # survivor count is exact
    np.testing.assert_allclose(got[1:5], ref[1:5], rtol=1e-12)  # sum w, sum w^2, sum w y, sum w y^2: fp64 throughout
    n = max(ref[1], 1.0)
    Gr, Gg = unpack(ref, d), unpack(got, d)
    mean = ref[5:5 + d] / n
    sd = np.sqrt(np.maximum(np.diag(Gr) / n - mean * mean, 0.0))
    sd = np.where(sd > 0, sd, np.maximum(np.abs(mean), 1.0))
    y_mean = ref[3] / n
    y_sd = np.sqrt(max(ref[4] / n - y_mean * y_mean, 0.0)) or max(abs(y_mean), 1.0)
    assert np.max(np.abs(Gg - Gr) / (n * np.outer(sd, sd))) < COV_TOL
    assert np.max(np.abs(got[5:5 + d] - ref[5:5 + d]) / (n * sd)) < 2e-7                     # sum w x: fp32 per 8 rows, fp64 across
    assert np.max(np.abs(got[5 + d:5 + 2 * d] - ref[5 + d:5 + 2 * d]) / (n * sd * y_sd)) < COV_TOL  # sum w x y


@pytest.mark.parametrize("d,n", [(1, 1), (2, 15), (9, 16), (9, 17), (16, 4096), (64, 50_001), (128, 5_000), (129, 5_003),
                                 (200, 12_345), (256, 20_011), (256, 300_000)])
@pytest.mark.parametrize("fused", [False, True])
def test_tf32_moments_vs_oracle(pkg, tf32, d, n, fused):
    from sparkdq4ml_b200 import ml
    guest, price, feats = table(n, d)
    dq = (guest, price) if fused else None
    ref = passes.wls_moments(feats, price, dq=dq, threads=8)
    dfeats, dprice = to_dev(*feats), to_dev(price)[0]
    got = ml.wls_moments(tf32, dfeats, dprice, dq=(dfeats[0], dprice) if fused else None)
    check_moments(got, ref, d)
    again = ml.wls_moments(tf32, dfeats, dprice, dq=(dfeats[0], dprice) if fused else None)
    assert again.tobytes() == got.tobytes()  # fixed-order folds everywhere: bit-reproducible
    host = ml.wls_moments(tf32, feats, price, dq=dq)  # host columns through the same entry point
    assert host.tobytes() == got.tobytes()


def test_tf32_spans_many_drains(pkg, tf32):
    """4 M rows: ~13 accumulation spans per CTA; the error must not grow with the number of spans."""
    from sparkdq4ml_b200 import ml
    d, n = 32, 4_000_000
    guest, price, feats = table(n, d)
    ref = passes.wls_moments(feats, price, dq=(guest, price), threads=16)
    dfeats, dprice = to_dev(*feats), to_dev(price)[0]
    check_moments(ml.wls_moments(tf32, dfeats, dprice, dq=(dfeats[0], dprice)), ref, d)


@pytest.mark.parametrize("d,n,params", [
    (16, 60_000, dict(reg_param=0.0)),
    (40, 30_000, dict(reg_param=0.1)),
    (256, 120_000, dict(reg_param=0.0)),
    (256, 120_000, dict(reg_param=0.3, elastic_net_param=0.5)),
    (130, 40_003, dict(reg_param=0.05, standardization=False)),
])
@pytest.mark.parametrize("weighted", [False, True])
def test_tf32_fit_vs_oracle(pkg, tf32, parity_log, d, n, params, weighted):
    guest, price, feats = table(n, d)
    base = dict(max_iter=100, reg_param=0.0, elastic_net_param=0.0, tol=1e-9, fit_intercept=True, standardization=True,
                solver="normal")
    base.update(params)
    p = spark_ml.LRParams(**base)
    w = None
    if weighted:
        rng = np.random.default_rng(11)
        w = rng.integers(0, 4, n).astype(np.float64) * rng.choice([0.5, 1.0, 1.75], n)
    ref = spark_ml.linear_regression_fit(feats, price, p, dq=(guest, price), weight=w)
    dfeats, dprice = to_dev(*feats), to_dev(price)[0]
    cols = {"guest": dfeats[0], "label": dprice}
    names = ["guest"]
    for k in range(1, d):
        cols[f"x{k}"] = dfeats[k]
        names.append(f"x{k}")
    if weighted:
        cols["w"] = to_dev(w)[0]
    df = pkg.VectorAssembler().setInputCols(names).transform(pkg.Frame(tf32, cols).with_fused_dq("guest", "label"))
    lr = pkg.LinearRegression().setMaxIter(p.max_iter).setRegParam(p.reg_param).setElasticNetParam(p.elastic_net_param)
    lr.setTol(p.tol).setStandardization(p.standardization).setSolver("normal")
    if weighted:
        lr.setWeightCol("w")
    model = lr.fit(df)
    # "1e-5 relative" per coefficient, with an explicit absolute floor: a coefficient smaller than COEF_FLOOR times the
    # largest one is compared on the floor's scale (the tf32 products of the Gram carry an absolute, not a relative, error)
    refc = np.asarray(ref.coefficients)
    diff = np.abs(model.coefficients() - refc)
    cmax = float(np.max(np.abs(refc)))
    per_coef = float(np.max(diff / np.maximum(np.abs(refc), COEF_FLOOR * cmax)))
    e_int = abs(model.intercept() - ref.intercept) / abs(ref.intercept)
    parity_log(f"tf32_fit_d{d}_n{n}_w{int(weighted)}_{params}", "per-coef rel (floor) / scale-rel / intercept rel",
               [per_coef, float(np.max(diff)) / cmax, e_int])
    assert float(np.max(diff)) / cmax < FIT_TOL and e_int < FIT_TOL and per_coef < PER_COEF_TOL
    assert model.num_rows == int(spark_ml._dq_keep_mask((guest, price), n).sum())


def test_tf32_nan_poisons_like_fp64(pkg, tf32):
    """A NaN in a surviving row must reach the moments of its column; a NaN in a dropped row must not."""
    from sparkdq4ml_b200 import ml
    n, d = 5_000, 12
    guest, price, feats = table(n, d)
    keep = spark_ml._dq_keep_mask((guest, price), n)
    dropped, kept = int(np.flatnonzero(~keep)[0]), int(np.flatnonzero(keep)[7])
    feats[3] = feats[3].copy()
    feats[3][dropped] = np.nan
    feats[5] = feats[5].copy()
    feats[5][dropped] = np.inf
    got = ml.wls_moments(tf32, feats, price, dq=(guest, price))
    assert np.all(np.isfinite(got))
    feats[7] = feats[7].copy()
    feats[7][kept] = np.nan
    got = ml.wls_moments(tf32, feats, price, dq=(guest, price))
    G = unpack(got, d)
    assert np.isnan(got[5 + 7]) and np.all(np.isnan(G[7, :])) and np.isfinite(G[3, 5]) and np.isfinite(got[5 + 3])


def test_tf32_negative_weight_is_an_error(pkg, tf32):
    guest, price, feats = table(3_000, 10)
    w = np.ones(3_000)
    w[77] = -2.0
    dfeats, dprice, dw = to_dev(*feats), to_dev(price)[0], to_dev(w)[0]
    cols = {"guest": dfeats[0], "label": dprice, "w": dw}
    names = ["guest"]
    for k in range(1, 10):
        cols[f"x{k}"] = dfeats[k]
        names.append(f"x{k}")
    df = pkg.VectorAssembler().setInputCols(names).transform(pkg.Frame(tf32, cols))
    with pytest.raises(pkg.capi.Dq4mlError) as ei:
        pkg.LinearRegression().setSolver("normal").setWeightCol("w").fit(df)
    assert ei.value.status == pkg.capi.ERR_INVALID


def test_tf32_wider_than_256_uses_the_fp64_kernel(pkg, ctx):
    from sparkdq4ml_b200 import ml
    n, d = 4_000, 260
    guest, price, feats = table(n, d)
    dfeats, dprice = to_dev(*feats), to_dev(price)[0]
    ref = ml.wls_moments(ctx, dfeats, dprice)
    ctx.set_gram_tier("tf32x3")
    try:
        got = ml.wls_moments(ctx, dfeats, dprice)
    finally:
        ctx.set_gram_tier("fp64")
    assert got.tobytes() == ref.tobytes()


def test_unknown_tier_is_refused(pkg, ctx):
    with pytest.raises(ValueError):
        ctx.set_gram_tier("bf16")
    assert pkg.capi.lib().dq4ml_ctx_set_gram_tier(ctx.handle, 7) == pkg.capi.ERR_INVALID


@pytest.mark.parametrize("d,n", [(256, 200_003), (40, 50_000), (3, 4_099)])
def test_tf32_tma_staging_equals_cp_async_staging(pkg, tf32, monkeypatch, d, n):
    """The tcgen05 kernel can stage a uniformly strided fp64 table (dq4ml_synth_generate's layout) with TMA tensor-map
    loads -- one 16-row x 256-column box per request, DQ4ML_TF32_TMA=1 -- instead of per-thread cp.async copies.  Only
    the staging differs, so the two paths must give bit-identical moments; a table whose columns are NOT one strided
    array must quietly take the cp.async path under the same setting."""
    from sparkdq4ml_b200 import ml
    g, p, x = tf32.synth_generate(n, d=d, p_low=0.05, p_high=0.05)
    strided = [g] + [x[k] for k in range(d - 1)]
    monkeypatch.setenv("DQ4ML_TF32_TMA", "0")
    ref = ml.wls_moments(tf32, strided, p, dq=(g, p))
    monkeypatch.setenv("DQ4ML_TF32_TMA", "1")
    got = ml.wls_moments(tf32, strided, p, dq=(g, p))
    assert got.tobytes() == ref.tobytes()
    scattered = [g] + [torch.empty(n + 16 * (1 + k % 3), dtype=torch.float64, device="cuda")[:n].copy_(x[k]) for k in range(d - 1)]
    torch.cuda.synchronize()
    assert ml.wls_moments(tf32, scattered, p, dq=(g, p)).tobytes() == ref.tobytes()
