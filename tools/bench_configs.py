#!/usr/bin/env python
"""Measurements for the BASELINE.json configs that are not the bench.py headline (configs 3, 4, 5) on ONE B200.

    python tools/bench_configs.py --config 3 [--rows 1000000000]    # 16 features, l-bfgs solver, elasticNet 0.5 (OWL-QN)
    python tools/bench_configs.py --config 4 [--rows 62500000]      # 256 features, normal equations (Gram)
    python tools/bench_configs.py --config 5                        # DQ-only filter sweep 1e7 .. 4e9 rows

Prints one JSON line per measurement (kept under profiles/).  Kernel times are CUDA events on the context's stream
(dq4ml_ctx_enable_pass_timing); inputs are far larger than the 126 MB L2.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)


def peak_hbm():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def frame(pkg, ctx, g, p, x, d):
    cols = {"guest": g, "label": p}
    names = ["guest"]
    for k in range(d - 1):
        cols[f"x{k + 1}"] = x[k]
        names.append(f"x{k + 1}")
    df = pkg.Frame(ctx, cols).with_fused_dq("guest", "label")
    return pkg.VectorAssembler().setInputCols(names).setOutputCol("features").transform(df)


def config3(pkg, ctx, rows, reps, row0=0, world=1, global_rows=None, d=16):
    import torch
    g, p, x = ctx.synth_generate(rows, d=d, row0=row0)
    df = frame(pkg, ctx, g, p, x, d)
    lr = pkg.LinearRegression().setMaxIter(40).setRegParam(1.0).setElasticNetParam(0.5).setSolver("l-bfgs")
    bytes_per_row = 4 + 8 * d  # int32 guest + 15 f64 features + f64 label(price); DQ columns are guest and label
    model = lr.fit(df)  # warm-up
    for _ in range(reps):
        ctx.enable_pass_timing(True)
        t0 = time.perf_counter()
        model = lr.fit(df)
        wall = time.perf_counter() - t0
        ms, passes_n = ctx.pass_timing()
        ctx.enable_pass_timing(False)
        gbs = bytes_per_row * rows * passes_n / (ms / 1e3) / 1e9
        if ctx.rank != 0:
            continue
        print(json.dumps({"config": 3, "n_gpus": world, "global_rows": global_rows or rows, "rows": rows, "features": d, "solver": model.solver_used,
                          "total_iterations": model.summary().totalIterations(), "data_passes": model.data_passes,
                          "fit_seconds": wall, "rows_per_sec_fit": (global_rows or rows) / wall, "pass_kernel_ms_avg": ms / passes_n,
                          "bytes_per_row_per_pass": bytes_per_row, "pass_gbs": gbs, "pass_frac_of_measured_hbm": gbs / peak_hbm(),
                          "coefficients_head": model.coefficients()[:4].tolist(), "intercept": model.intercept()}), flush=True)
    del g, p, x
    torch.cuda.empty_cache()


def config4(pkg, ctx, rows, reps, row0=0, world=1, tier="fp64"):
    import torch
    from sparkdq4ml_b200 import ml
    d = 256
    ctx.set_gram_tier(tier)
    g, p, x = ctx.synth_generate(rows, d=d, row0=row0)
    feats = [g] + [x[k] for k in range(d - 1)]
    flop_per_row = d * (d + 1) + 4 * d  # symmetric half of X^T X at 2 flop/MAC + X^T y + sums
    bytes_per_row = 4 + 8 * d
    ml.wls_moments(ctx, feats, p, dq=(g, p))
    for _ in range(reps):
        ctx.enable_pass_timing(True)
        t0 = time.perf_counter()
        mom = ml.wls_moments(ctx, feats, p, dq=(g, p))
        wall = time.perf_counter() - t0
        ms, passes_n = ctx.pass_timing()
        ctx.enable_pass_timing(False)
        if ctx.rank != 0:
            continue
        print(json.dumps({"config": 4, "gram_tier": tier, "n_gpus": world, "rows_per_gpu": rows, "features": d, "survivors_all_ranks": mom[0], "gram_pass_ms": ms / passes_n,
                          "wall_ms": wall * 1e3, "rows_per_sec_pass": rows / (ms / passes_n / 1e3),
                          "algorithmic_tflops": flop_per_row * rows / (ms / passes_n / 1e3) / 1e12,
                          "frac_of_measured_fp64_peak_37.2": flop_per_row * rows / (ms / passes_n / 1e3) / 1e12 / 37.2,
                          "gbs": bytes_per_row * rows / (ms / passes_n / 1e3) / 1e9,
                          "frac_of_measured_hbm": bytes_per_row * rows / (ms / passes_n / 1e3) / 1e9 / peak_hbm()}), flush=True)
    df = frame(pkg, ctx, g, p, x, d)
    lr = pkg.LinearRegression().setMaxIter(40).setRegParam(0.0).setSolver("normal")
    t0 = time.perf_counter()
    model = lr.fit(df)
    wall = time.perf_counter() - t0
    if ctx.rank == 0:
        print(json.dumps({"config": 4, "gram_tier": tier, "n_gpus": world, "what": "whole fit (Gram pass + all-reduce + packed Cholesky on the host)", "rows_per_gpu": rows,
                      "fit_seconds": wall, "rows_per_sec_fit": rows * world / wall, "solver": model.solver_used,
                      "coefficients_head": model.coefficients()[:4].tolist(), "intercept": model.intercept()}), flush=True)
    del g, p, x, feats
    torch.cuda.empty_cache()


def config5(pkg, ctx, sizes, reps):
    import torch
    stream = torch.cuda.ExternalStream(ctx.stream)
    for n in sizes:
        g, p, _ = ctx.synth_generate(n, d=1)
        go, po, n1 = pkg.dq_filter(ctx, g, p)  # warm-up (also sizes the tile-status scratch)
        m = go.shape[0]
        best = 1e30
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            del go, po  # the outputs are as large as the inputs: free the previous pair before allocating the next
            ctx.synchronize()
            e0.record(stream)
            go, po, n1 = pkg.dq_filter(ctx, g, p)
            e1.record(stream)
            ctx.synchronize()
            best = min(best, e0.elapsed_time(e1))
        bytes_alg = 12 * n + 12 * m
        print(json.dumps({"config": 5, "rows": n, "survivors": m, "after_rule1": n1, "call_ms_best": best,
                          "rows_per_sec": n / (best / 1e3), "algorithmic_bytes": bytes_alg,
                          "gbs": bytes_alg / (best / 1e3) / 1e9, "frac_of_measured_hbm": bytes_alg / (best / 1e3) / 1e9 / peak_hbm(),
                          "digest": [str(v) for v in ctx.rows_digest(go, po, seed=5)],
                          "note": "whole dq4ml_dq_filter call: 2 memsets + kernel + 16 B D2H + sync"}), flush=True)
        del g, p, go, po
        torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True)
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--features", type=int, default=16, help="config 3 only: number of feature columns (the config itself names 16)")
    ap.add_argument("--gram-tier", type=str, default="fp64", choices=["fp64", "tf32x3"])
    ap.add_argument("--sizes", type=str, default="10000000,100000000,1000000000,4000000000")
    a = ap.parse_args()
    import sparkdq4ml_b200 as pkg
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:  # torchrun: config 3 is strong-scaled (global rows sharded), config 4 weak-scaled (rows per GPU)
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    ctx = pkg.Context(local).bind_torch_distributed()
    if a.config == 3:
        n_global = a.rows or 1_000_000_000
        lo, hi = pkg.shard_bounds(n_global, rank, world)
        config3(pkg, ctx, hi - lo, a.reps, row0=lo, world=world, global_rows=n_global, d=a.features)
    elif a.config == 4:
        rows = a.rows or 62_500_000
        config4(pkg, ctx, rows, a.reps, row0=rank * rows, world=world, tier=a.gram_tier)
    elif a.config == 5:
        config5(pkg, ctx, [int(s) for s in a.sizes.split(",")], a.reps)
    ctx.close()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
