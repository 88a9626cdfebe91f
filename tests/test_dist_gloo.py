"""CPU, world_size 2 over gloo: the row-sharded protocol.  Each rank aggregates its own shard (here with the oracle,
standing in for the GPU pass), the result vectors are summed across ranks, and every rank runs the product's host
solver on the sum -- the result must equal the single-process fit."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import sparkdq4ml_b200 as pkg
    from sparkdq4ml_b200 import ml
    from sparkdq4ml_b200.runtime import torch_allreduce_callback
    from oracle import passes, spark_ml

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        d = 3
        lo, hi = pkg.shard_bounds(n, rank, world)
        guest, price, extra = passes.synth_generate(hi - lo, d=d, row0=lo, threads=1)
        feats = [guest, extra[0], extra[1]]
        # shard-local aggregation, then the cross-rank sum the library performs after every pass
        mom = passes.wls_moments(feats, price, dq=(guest, price))
        torch_allreduce_callback()(mom)
        lr = pkg.LinearRegression().setMaxIter(40).setRegParam(0.3).setElasticNetParam(0.5)
        model = ml.wls_solve(mom, d, lr)
        # l-bfgs branch: summarizer state is merged as (count, mean, M2n) in rank order; loss/grad sums are added
        s = passes.summarizer(feats, price, dq=(guest, price))
        cnt = float(s["count"])
        local = np.concatenate([[cnt], s["mean"] * cnt, s["variance"] * (cnt - 1) + cnt * s["mean"] ** 2])
        torch_allreduce_callback()(local)
        tot = local[0]
        mean = local[1:1 + d + 1] / tot
        var = (local[1 + d + 1:] - tot * mean ** 2) / (tot - 1)
        q.put((rank, model.coefficients().tolist(), model.intercept(), mom.tolist(), mean.tolist(), var.tolist(), lo, hi))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_sharded_wls_equals_single_process():
    import sparkdq4ml_b200 as pkg
    from sparkdq4ml_b200 import ml
    from oracle import passes

    n, world, d = 40_000, 2, 3
    port = _free_port()
    ctxmp = mp.get_context("spawn")
    q = ctxmp.Queue()
    procs = [ctxmp.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=240) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert results[0][6] == 0 and results[0][7] == results[1][6] and results[1][7] == n
    guest, price, extra = passes.synth_generate(n, d=d, threads=1)
    feats = [guest, extra[0], extra[1]]
    mom = passes.wls_moments(feats, price, dq=(guest, price))
    lr = pkg.LinearRegression().setMaxIter(40).setRegParam(0.3).setElasticNetParam(0.5)
    single = ml.wls_solve(mom, d, lr)
    s = passes.summarizer(feats, price, dq=(guest, price))
    for r in results:
        assert r[3][0] == mom[0]  # survivor count is exact
        np.testing.assert_allclose(r[3], mom, rtol=1e-12)
        np.testing.assert_allclose(r[1], single.coefficients(), rtol=1e-9)
        np.testing.assert_allclose(r[2], single.intercept(), rtol=1e-9)
        np.testing.assert_allclose(r[4], s["mean"], rtol=1e-12)
        np.testing.assert_allclose(r[5], s["variance"], rtol=1e-9)
    assert results[0][1] == results[1][1]  # every rank lands on bit-identical coefficients
