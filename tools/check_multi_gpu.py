#!/usr/bin/env python
"""Run under torchrun on N >= 2 GPUs: the row-sharded fit (one process per GPU, NCCL all-reduce of the result vector
inside libdq4ml) must equal the single-GPU fit over the same global rows.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/check_multi_gpu.py
"""
import faulthandler
import json
import os
import sys

faulthandler.dump_traceback_later(300, exit=True)

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import sparkdq4ml_b200 as pkg  # noqa: E402


def frame(ctx, g, p, x, d, label="label", weighted=False):
    cols = {"guest": g, label: p}
    if weighted:  # a weight that depends on the row's own data only, so that every sharding sees the same column
        cols["w"] = (g % 4).double() * 0.5
    names = ["guest"]
    for k in range(d - 1):
        cols[f"x{k + 1}"] = x[k]
        names.append(f"x{k + 1}")
    df = pkg.Frame(ctx, cols).with_fused_dq("guest", label)
    return pkg.VectorAssembler().setInputCols(names).transform(df)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    results = {}
    n_global = int(os.environ.get("DQ4ML_CHECK_ROWS", 8_000_000)) + 4321
    import time
    t0 = time.time()
    cases = [("wls_d1", 1, dict(reg=1.0, alpha=1.0, solver="auto")),
             ("owlqn_d16", 16, dict(reg=0.3, alpha=0.5, solver="l-bfgs")),
             ("owlqn_d24", 24, dict(reg=0.3, alpha=0.5, solver="l-bfgs")),
             ("gram_d40", 40, dict(reg=0.1, alpha=0.0, solver="normal")),
             ("gram_d40_weighted", 40, dict(reg=0.1, alpha=0.0, solver="normal", weighted=True)),
             ("lbfgs_d12_weighted", 12, dict(reg=0.2, alpha=0.0, solver="l-bfgs", weighted=True)),
             ("gram_d200_tf32x3", 200, dict(reg=0.0, alpha=0.0, solver="normal", tier="tf32x3", rtol=1e-5))]
    failures = []
    sharded = pkg.Context(local).bind_torch_distributed()
    single = pkg.Context(local)  # unbound: sees only what it is given
    for name, d, hp in cases:
        lo, hi = pkg.shard_bounds(n_global, rank, world)
        g, p, x = sharded.synth_generate(hi - lo, d=d, row0=lo)
        lr = pkg.LinearRegression().setMaxIter(40).setRegParam(hp["reg"]).setElasticNetParam(hp["alpha"]).setSolver(hp["solver"])
        wt = hp.get("weighted", False)
        if wt:
            lr.setWeightCol("w")
        sharded.set_gram_tier(hp.get("tier", "fp64"))
        single.set_gram_tier(hp.get("tier", "fp64"))
        rtol = hp.get("rtol", 1e-8)  # the tcgen05 tier sums fp32 partial results in a sharding-dependent grouping
        m = lr.fit(frame(sharded, g, p, x, d, weighted=wt))
        ev = m.evaluate(frame(sharded, g, p, x, d, weighted=wt))
        res = {"coef": m.coefficients().tolist(), "intercept": m.intercept(), "rows": m.num_rows, "rmse": ev.rootMeanSquaredError(),
               "hist": m.summary().objectiveHistory().tolist()}
        # every rank must hold the identical model
        blob = [None] * world
        dist.all_gather_object(blob, res)
        assert all(b == blob[0] for b in blob), f"{name}: ranks disagree"
        if rank == 0:
            gg, pp, xx = single.synth_generate(n_global, d=d, row0=0)
            ref = lr.fit(frame(single, gg, pp, xx, d, weighted=wt))
            ref_ev = ref.evaluate(frame(single, gg, pp, xx, d, weighted=wt))
            try:  # a failed comparison must not keep rank 0 from the barrier the other ranks wait in
                assert ref.num_rows == res["rows"], (ref.num_rows, res["rows"])
                scale = max(1.0, float(np.max(np.abs(ref.coefficients()))))  # the tolerance of a tier is relative to the largest coefficient
                np.testing.assert_allclose(res["coef"], ref.coefficients(), rtol=rtol, atol=max(1e-11, rtol * scale if "tier" in hp else 0.0))
                np.testing.assert_allclose(res["intercept"], ref.intercept(), rtol=max(rtol, 1e-9))
                np.testing.assert_allclose(res["rmse"], ref_ev.rootMeanSquaredError(), rtol=max(rtol, 1e-9))
                np.testing.assert_allclose(res["hist"], ref.summary().objectiveHistory(), rtol=rtol)
            except AssertionError as e:
                failures.append(f"{name}: {str(e)[:400]}")
            results[name] = {"rows": res["rows"], "intercept": res["intercept"], "iterations": len(res["hist"])}
            print(f"[check] {name} ok after {time.time() - t0:.1f} s", file=sys.stderr, flush=True)
        dist.barrier()
    if rank == 0:
        print(json.dumps({"multi_gpu_check": "ok" if not failures else "FAILED", "world": world, "cases": results, "failures": failures}))
    dist.destroy_process_group()
    if failures:
        sys.exit(1)


if __name__ == "__main__":
    main()
