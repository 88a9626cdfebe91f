"""GPU parity, DQ section: the UDF column kernels and the fused filter-compaction against the oracle, through the C ABI."""
import hashlib

import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import passes

pytestmark = pytest.mark.gpu


def edge_rows():
    nan, inf = float("nan"), float("inf")
    prices = [nan, inf, -inf, 0.0, -0.0, 19.999999999999996, 20.0, 20.000000000000004, 89.99, 90.0, 90.00000000000001,
              5e-324, -1.0, 1e308, 150.0, 23.24]
    guests = [13, 14, 1, 35, 0, -5, 2**31 - 1, -2**31]
    g, p = np.meshgrid(np.asarray(guests, dtype=np.int32), np.asarray(prices, dtype=np.float64), indexing="ij")
    return np.ascontiguousarray(g.ravel()), np.ascontiguousarray(p.ravel())


@pytest.mark.parametrize("where", ["device", "host"])
def test_udf_columns_bit_exact(pkg, ctx, where):
    g, p = edge_rows()
    rng = np.random.default_rng(1)
    g = np.concatenate([g, rng.integers(1, 36, 10_001).astype(np.int32)])
    p = np.concatenate([p, rng.uniform(0, 160, 10_001)])
    u1, u2 = pkg.MinimumPriceDataQualityUdf(), pkg.PriceCorrelationDataQualityUdf()
    if where == "device":
        gd, pd = torch.from_numpy(g).cuda(), torch.from_numpy(p).cuda()
        o1 = u1.call_columns(ctx, pd).cpu().numpy()
        o2 = u2.call_columns(ctx, pd, gd).cpu().numpy()
    else:
        o1, o2 = u1.call_columns(ctx, p), u2.call_columns(ctx, p, g)
    assert o1.tobytes() == passes.udf_minimum_price(p).tobytes()
    assert o2.tobytes() == passes.udf_price_correlation(p, g).tobytes()


def test_udf_null_semantics(pkg, ctx):
    p = np.array([50.0, 95.0, 10.0, 60.0, 70.0])
    g = np.array([20, 5, 20, 20, 20], dtype=np.int32)
    pv = np.array([1, 1, 1, 0, 1], dtype=np.uint8)
    gv = np.array([1, 1, 1, 1, 0], dtype=np.uint8)
    u1, u2 = pkg.MinimumPriceDataQualityUdf(), pkg.PriceCorrelationDataQualityUdf()
    out = u2.call_columns(ctx, p, g, price_valid=pv, guest_valid=gv)
    assert out.tolist() == passes.udf_price_correlation(p, g, pv, gv).tolist() == [50.0, -1.0, 10.0, -1.0, -1.0]
    with pytest.raises(pkg.capi.NullValueError):  # MinUdf:12 -> NullPointerException
        u1.call_columns(ctx, p, price_valid=pv)
    assert u1.call_columns(ctx, p, price_valid=np.ones(5, dtype=np.uint8)).tolist() == [50.0, 95.0, -1.0, 60.0, 70.0]


@pytest.mark.parametrize("name", ["abstract", "small", "full"])
def test_filter_on_reference_fixtures(pkg, ctx, golden, name):
    doc = golden("dataset_" + name)
    g = np.asarray(doc["guest"], dtype=np.int32)
    p = np.asarray(doc["price"], dtype=np.float64)
    go, po, n1 = pkg.dq_filter(ctx, torch.from_numpy(g).cuda(), torch.from_numpy(p).cuda())
    go, po = go.cpu().numpy(), po.cpu().numpy()
    assert (n1, len(go)) == (doc["n_after_rule1"], doc["n_after_rule2"])
    packed = np.zeros(len(go), dtype=[("g", "<i4"), ("p", "<f8")])
    packed["g"], packed["p"] = go, po
    assert hashlib.sha256(packed.tobytes()).hexdigest() == doc["survivor_sha256"]
    # host buffers through the same entry point
    gh, ph, n1h = pkg.dq_filter(ctx, g, p)
    assert gh.tobytes() == go.tobytes() and ph.tobytes() == po.tobytes() and n1h == n1


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 2047, 2048, 2049, 65_537, 1_000_003])
@pytest.mark.parametrize("rates", [(0.01, 0.01), (0.25, 0.25)])
def test_filter_bit_exact_vs_oracle(pkg, ctx, n, rates):
    guest, price, _ = passes.synth_generate(n, d=1, p_low=rates[0], p_high=rates[1])
    eg, ep = edge_rows()
    if n >= eg.size:  # sprinkle NaN/Inf/boundary rows in
        guest[:eg.size], price[:eg.size] = eg, ep
    ref_g, ref_p, ref_n1 = passes.dq_filter(guest, price)
    go, po, n1 = pkg.dq_filter(ctx, torch.from_numpy(guest).cuda(), torch.from_numpy(price).cuda())
    assert n1 == ref_n1 and go.shape[0] == ref_g.size
    assert go.cpu().numpy().tobytes() == ref_g.tobytes()
    assert po.cpu().numpy().tobytes() == ref_p.tobytes()


def test_filter_stage_params_reproduce_two_step_flow(pkg, ctx):
    """rule 1 + filter, then rule 2 + filter as two calls (the lab's App:68-78 then App:86-90) == the fused call."""
    from sparkdq4ml_b200 import dq as dqmod
    guest, price, _ = passes.synth_generate(100_000, d=1, p_low=0.1, p_high=0.2)
    g1, p1, _ = pkg.dq_filter(ctx, guest, price, dqmod.rule1_only_params())
    r1 = passes.udf_minimum_price(price)
    assert g1.tobytes() == guest[r1 > 0].tobytes() and p1.tobytes() == price[r1 > 0].tobytes()
    g2, p2, _ = pkg.dq_filter(ctx, g1, p1, dqmod.rule2_only_params())
    gf, pf, n1 = pkg.dq_filter(ctx, guest, price)
    assert g2.tobytes() == gf.tobytes() and p2.tobytes() == pf.tobytes() and n1 == g1.size


def test_filter_rejects_bad_arguments(pkg, ctx):
    import ctypes as C
    g = torch.zeros(8, dtype=torch.int32, device="cuda")
    p = torch.zeros(8, dtype=torch.float64, device="cuda")
    n_out = C.c_int64(0)
    L = pkg.capi.lib()
    rc = L.dq4ml_dq_filter(ctx.handle, g.data_ptr(), p.data_ptr(), 8, None, g.data_ptr(), p.data_ptr(), C.byref(n_out), None)
    assert rc == pkg.capi.ERR_INVALID  # in-place
    go, po = torch.zeros_like(g), torch.zeros_like(p)
    rc = L.dq4ml_dq_filter(ctx.handle, g.data_ptr() + 4, p.data_ptr(), 4, None, go.data_ptr(), po.data_ptr(), C.byref(n_out), None)
    assert rc == pkg.capi.ERR_INVALID and b"aligned" in L.dq4ml_last_error()


def test_generator_and_digest_bit_identical_to_oracle(pkg, ctx, golden):
    doc = golden("synth_small")
    n, d = 300_001, 5
    g, p, x = ctx.synth_generate(n, d=d, row0=12345, p_low=0.03, p_high=0.02)
    og, op, ox = passes.synth_generate(n, d=d, row0=12345, p_low=0.03, p_high=0.02)
    assert g.cpu().numpy().tobytes() == og.tobytes()
    assert p.cpu().numpy().tobytes() == op.tobytes()
    assert x.cpu().numpy().tobytes() == ox.tobytes()
    assert ctx.rows_digest(g, p, seed=9) == passes.rows_digest(og, op, seed=9)
    g4, p4, _ = ctx.synth_generate(doc["n"], d=doc["d"], p_low=doc["p_low"], p_high=doc["p_high"])
    assert [str(v) for v in ctx.rows_digest(g4, p4)] == doc["digest"]


@pytest.mark.parametrize("n", [10_000_000, 100_000_000])
def test_filter_at_scale_by_digest_and_properties(pkg, ctx, n):
    """BASELINE config 5 sizes: bit-exactness through the order-sensitive digest + size-independent properties."""
    g, p, _ = ctx.synth_generate(n, d=1)
    go, po, n1 = pkg.dq_filter(ctx, g, p)
    m = go.shape[0]
    og, op, _ = passes.synth_generate(n, d=1)
    ref_g, ref_p, ref_n1 = passes.dq_filter(og, op)
    assert (m, n1) == (ref_g.size, ref_n1)
    assert ctx.rows_digest(go, po, seed=3) == passes.rows_digest(ref_g, ref_p, seed=3)
    # idempotence: filtering the survivors again drops nothing
    g2, p2, n1b = pkg.dq_filter(ctx, go, po)
    assert g2.shape[0] == m and n1b == m
    assert ctx.rows_digest(g2, p2, seed=3) == ctx.rows_digest(go, po, seed=3)
    # every survivor satisfies both rules
    assert bool((po >= 20).all()) and not bool(((go < 14) & (po > 90)).any())
