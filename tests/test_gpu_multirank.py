"""GPU: the row-sharded (multi-rank) path of the library itself -- every rank runs the sm_100a passes over its own shard
and the library sums the result vectors across ranks -- against the single-process fit and the oracle.

Two transports are covered:
  * the host-callback all-reduce (gloo), with BOTH ranks sharing cuda:0, so the cross-rank logic of the library
    (per-rank min/max slots, shared summarizer centres, empty shards, rank-invariant kernel choice) is exercised on a
    one-GPU box too;
  * the in-stream NCCL all-reduce / peer-mailbox epilogue, one rank per GPU; skipped when fewer than 2 GPUs are visible.

Also here: BASELINE config 3's exact hyper-parameters (l-bfgs solver, regParam 1, elasticNet 0.5, 16 features) at
n = 10^7 rows against the oracle (Spark's treeAggregate replaced by Spark's Aggregator.merge = plain sums).
"""
import hashlib
import os
import socket
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT, rel_err, scale_err

pytestmark = pytest.mark.gpu

CONFIG3 = dict(max_iter=40, reg_param=1.0, elastic_net_param=0.5, solver="l-bfgs")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _frame(pkg, ctx, g, p, x, d):
    cols = {"guest": g, "label": p}
    names = ["guest"]
    for k in range(d - 1):
        cols[f"x{k + 1}"] = x[k]
        names.append(f"x{k + 1}")
    df = pkg.Frame(ctx, cols).with_fused_dq("guest", "label")
    return pkg.VectorAssembler().setInputCols(names).setOutputCol("features").transform(df)


def _estimator(pkg, p):
    lr = pkg.LinearRegression().setMaxIter(p["max_iter"]).setRegParam(p["reg_param"])
    return lr.setElasticNetParam(p["elastic_net_param"]).setSolver(p.get("solver", "auto"))


def _same(a, b):
    return repr(a) == repr(b)  # NaN-aware equality of nested plain values


def _digest(g, p):
    return hashlib.sha256(g.cpu().numpy().tobytes() + p.cpu().numpy().tobytes()).hexdigest()


def _run_all(pkg, ctx, n, lo, hi, nan_row=None, keep_arrays=False):
    """Everything a rank computes over rows [lo, hi) of the n-row synthetic table; the cross-rank sums happen inside the
    library.  Returns plain Python values."""
    from sparkdq4ml_b200 import ml
    out = {}
    d = 16
    g, p, x = ctx.synth_generate(hi - lo, d=d, row0=lo, p_low=0.03, p_high=0.03)
    if nan_row is not None and lo <= nan_row < hi:
        p[nan_row - lo] = float("nan")  # a NaN price survives both rules (Spark SQL ordering) and reaches the statistics
    df = _frame(pkg, ctx, g, p, x, d)
    if nan_row is None:
        m = _estimator(pkg, CONFIG3).fit(df)
        out["c3"] = (m.coefficients().tolist(), m.intercept(), m.summary().objectiveHistory().tolist(), m.data_passes,
                     m.num_rows, m.solver_used)
        # normal equations, 3 features (streaming WLS kernel) and 12 features (Gram kernel)
        for dd, key in ((3, "wls3"), (12, "wls12")):
            dfd = _frame(pkg, ctx, g, p, x, dd)
            mm = _estimator(pkg, dict(max_iter=40, reg_param=0.3, elastic_net_param=0.5)).fit(dfd)
            out[key] = (mm.coefficients().tolist(), mm.intercept(), mm.num_rows)
    # evaluation of a fixed model (a NaN label turns a *fitted* model into NaNs, which would hide the min/max semantics)
    fixed = pkg.LinearRegressionModel.from_coefficients([5.0] + [0.25 * ((k % 5) - 2) for k in range(1, d)], 22.0)
    ev = fixed.evaluate(df)
    out["eval"] = (ev.numInstances(), ev.rootMeanSquaredError(), ev.r2(), ev.meanAbsoluteError(), ev.devianceResiduals())
    feats = [g] + [x[k] for k in range(3)]
    s = ml.summarize(ctx, feats, p, dq=(g, p))
    out["summ"] = {k: (v.tolist() if hasattr(v, "tolist") else v) for k, v in s.items()}
    # the DQ section itself, row-sharded: local compaction + the global position of this shard's survivors
    go, po, n1 = pkg.dq_filter(ctx, g, p)
    off, tot = pkg.dq_filter_global_offset(ctx, go.shape[0])
    out["filter"] = (off, tot, int(go.shape[0]), n1, _digest(go, po))
    if keep_arrays:
        out["filter_arrays"] = (go, po)
    return out


def _worker(rank, world, port, backend, n, devices, q, nan_row, tier):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist

    import sparkdq4ml_b200 as pkg

    dev = devices[rank]
    torch.cuda.set_device(dev)
    kw = {"device_id": torch.device(f"cuda:{dev}")} if backend == "nccl" else {}
    dist.init_process_group(backend, init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world, **kw)
    try:
        ctx = pkg.Context(dev).bind_torch_distributed()
        ctx.set_gram_tier(tier)
        lo, hi = pkg.shard_bounds(n, rank, world)
        res = _run_all(pkg, ctx, n, lo, hi, nan_row)
        res["bounds"] = (lo, hi)
        q.put((rank, res))
        ctx.close()
    finally:
        dist.destroy_process_group()


def _spawn(world, backend, n, devices, nan_row=None, tier="fp64"):
    import torch.multiprocessing as mp
    port = _free_port()
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    procs = [mpc.Process(target=_worker, args=(r, world, port, backend, n, devices, q, nan_row, tier)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted((q.get(timeout=600) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    return [r for _, r in got]


def _check_against_single(results, single, n, parity_log, tag):
    assert results[0]["bounds"][0] == 0 and results[-1]["bounds"][1] == n
    for a, b in zip(results, results[1:]):
        assert a["bounds"][1] == b["bounds"][0]
    for r in results[1:]:  # the replicated host optimiser lands on bit-identical iterates on every rank
        for key in ("c3", "wls3", "wls12", "eval", "summ"):
            assert _same(r.get(key), results[0].get(key)), key
    # sharded filter: offsets are the exclusive scan of the shard counts, and shard r holds exactly the slice
    # [offset, offset + m) of the single-process output
    go, po = single["filter_arrays"]
    pos = 0
    for r in results:
        off, tot, m, n1, dig = r["filter"]
        assert off == pos and tot == single["filter"][2]
        assert dig == _digest(go[off:off + m], po[off:off + m])
        pos += m
    assert pos == single["filter"][2] and sum(r["filter"][3] for r in results) == single["filter"][3]
    r = results[0]
    if "c3" in single:
        assert r["c3"][3:] == single["c3"][3:]  # data passes, rows entering the fit, solver
        assert len(r["c3"][2]) == len(single["c3"][2])
        e_coef = scale_err(r["c3"][0], single["c3"][0])
        e_int = rel_err(r["c3"][1], single["c3"][1])
        e_hist = rel_err(r["c3"][2], single["c3"][2])
        parity_log(tag, "config3 N-rank vs 1-rank: coef (scale-rel) / intercept / objectiveHistory", [e_coef, e_int, e_hist])
        assert e_coef < 1e-9 and e_int < 1e-9 and e_hist < 1e-9
        for key in ("wls3", "wls12"):
            assert r[key][2] == single[key][2]
            assert scale_err(r[key][0], single[key][0]) < 1e-9 and rel_err(r[key][1], single[key][1]) < 1e-9
    assert r["eval"][0] == single["eval"][0]
    for i in (1, 2, 3):
        np.testing.assert_allclose(r["eval"][i], single["eval"][i], rtol=1e-9, equal_nan=True)
    np.testing.assert_allclose(r["eval"][4], single["eval"][4], rtol=1e-12, equal_nan=True)
    s, t = r["summ"], single["summ"]
    assert s["count"] == t["count"]
    for key in ("min", "max", "nnz"):
        assert np.array_equal(np.asarray(s[key]), np.asarray(t[key]), equal_nan=True), key  # exact, NaN-aware
    for key in ("mean", "variance", "normL1", "normL2"):
        np.testing.assert_allclose(s[key], t[key], rtol=1e-10, equal_nan=True, err_msg=key)


def test_config3_hyperparameters_at_1e7_rows_vs_oracle(pkg, ctx, parity_log):
    """BASELINE config 3 as benchmarked (l-bfgs solver => OWL-QN, regParam 1, elasticNet 0.5, 16 features, fused DQ) at
    n = 10^7 against the oracle: same number of optimiser states and data passes, coefficients / intercept / objective."""
    from oracle import passes, spark_ml
    n, d = 10_000_000, 16
    guest, price, extra = passes.synth_generate(n, d=d)
    feats = [guest] + [np.ascontiguousarray(extra[k]) for k in range(d - 1)]
    ref = spark_ml.linear_regression_fit(feats, price, spark_ml.LRParams(**CONFIG3), dq=(guest, price),
                                         threads=passes.host_threads())
    g, p, x = ctx.synth_generate(n, d=d)
    model = _estimator(pkg, CONFIG3).fit(_frame(pkg, ctx, g, p, x, d))
    assert model.solver_used == ref.solver_used == "owlqn"
    s = model.summary()
    assert s.totalIterations() == ref.total_iterations and model.data_passes == ref.data_passes
    e_hist = rel_err(s.objectiveHistory(), ref.objective_history)
    e_coef = scale_err(model.coefficients(), ref.coefficients)
    e_int = rel_err(model.intercept(), ref.intercept)
    parity_log("config3_1e7_vs_oracle", "objectiveHistory rel / coef scale-rel / intercept rel", [e_hist, e_coef, e_int])
    assert e_hist < 1e-9 and e_coef < 1e-9 and e_int < 1e-9
    # which coefficients the L1 term zeroes is part of the result
    assert (model.coefficients() == 0.0).tolist() == (np.asarray(ref.coefficients) == 0.0).tolist()


@pytest.mark.timeout(900)
@pytest.mark.parametrize("nan_row", [None, 1_500_123])
def test_two_ranks_host_callback_on_one_gpu(pkg, ctx, parity_log, nan_row):
    n = 2_000_000
    single = _run_all(pkg, ctx, n, 0, n, nan_row, keep_arrays=True)
    results = _spawn(2, "gloo", n, [0, 0], nan_row)
    _check_against_single(results, single, n, parity_log, f"gloo2_nan{nan_row}")
    if nan_row is not None:  # the NaN residual: SQL max() returns it, SQL min() does not
        lo, hi = single["eval"][4]
        assert np.isnan(hi) and np.isfinite(lo)
        # MultivariateOnlineSummarizer's min / max never pick a NaN up; its sums do
        assert np.isfinite(single["summ"]["min"]).all() and np.isfinite(single["summ"]["max"]).all()
        assert np.isnan(single["summ"]["mean"][-1])


@pytest.mark.timeout(900)
@pytest.mark.parametrize("tier", ["fp64", "tf32x3"])
def test_two_ranks_with_an_empty_shard(pkg, ctx, tier):
    """n < world x 1024 leaves rank 0 without rows: it must still take part in every cross-rank sum, with a vector of
    the same length as its peer's, whichever kernel tier is selected."""
    n = 1500
    assert pkg.shard_bounds(n, 0, 2) == (0, 0)
    ctx.set_gram_tier(tier)
    try:
        single = _run_all(pkg, ctx, n, 0, n)
    finally:
        ctx.set_gram_tier("fp64")
    results = _spawn(2, "gloo", n, [0, 0], None, tier)
    assert results[0]["bounds"] == (0, 0)
    for key in ("c3", "wls3", "wls12", "eval", "summ"):
        assert _same(results[0][key], results[1][key]), key
    assert results[0]["filter"][:3] == (0, single["filter"][2], 0) and results[1]["filter"][:3] == (0, single["filter"][2], single["filter"][2])
    tol = 1e-9 if tier == "fp64" else 1e-5
    assert results[0]["c3"][3:] == single["c3"][3:]
    assert scale_err(results[0]["c3"][0], single["c3"][0]) < 1e-9
    for key in ("wls3", "wls12"):
        assert results[0][key][2] == single[key][2]
        assert scale_err(results[0][key][0], single[key][0]) < tol


@pytest.mark.timeout(900)
def test_two_ranks_nccl_one_gpu_each(pkg, ctx, parity_log):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    n = 4_000_000
    single = _run_all(pkg, ctx, n, 0, n, keep_arrays=True)
    results = _spawn(2, "nccl", n, [0, 1])
    _check_against_single(results, single, n, parity_log, "nccl2")
