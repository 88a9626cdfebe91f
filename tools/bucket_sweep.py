"""Per-pass HBM fractions of the streaming passes of an l-bfgs fit and of the evaluation, by feature count.

    python tools/bucket_sweep.py --rows 200000000 --features 2,4,8,16,17,20,24,25,28,32 [--weighted]

For every d: one LinearRegression(solver="l-bfgs", maxIter=3) fit over a synthetic table (guest int32 + d-1 fp64 columns,
label = price, fused DQ) with per-pass timing on, then model.evaluate (wall clock, mean of 3).  Passes of a fit in launch order: the column heads
(a 1-CTA kernel, not reported), the summarizer, then one loss/gradient pass per distinct evaluation point.  A pass reads
4 + 8 (d - 1) bytes of features + 8 of label (+ 8 of weight) per row; the DQ columns are feature 0 and the label.
Prints one JSON line per d (fractions are of MEASURED_PEAKS.json hbm_gbs, else 6553).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparkdq4ml_b200 as pkg  # noqa: E402
from sparkdq4ml_b200 import ml  # noqa: E402


def peak_gbs():
    try:
        with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"])
    except Exception:
        return 6553.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=200_000_000)
    ap.add_argument("--features", type=str, default="2,4,8,16,17,20,24,25,28,32")
    ap.add_argument("--weighted", action="store_true")
    ap.add_argument("--loss", type=str, default="squaredError")
    args = ap.parse_args()
    peak = peak_gbs()
    ctx = pkg.Context(0)
    for d in [int(x) for x in args.features.split(",")]:
        rows = args.rows if d <= 32 else args.rows // 2
        g, p, x = ctx.synth_generate(rows, d=d)
        cols = {"guest": g, "label": p}
        names = ["guest"]
        for k in range(d - 1):
            cols["f%d" % k] = x[k]
            names.append("f%d" % k)
        if args.weighted:
            import torch
            cols["w"] = torch.rand(rows, dtype=torch.float64, device="cuda:0") + 0.5
        df = ml.Frame(ctx, cols).with_fused_dq("guest", "label")
        df = ml.VectorAssembler().setInputCols(names).setOutputCol("features").transform(df)
        lr = ml.LinearRegression().setMaxIter(3).setRegParam(0.1).setElasticNetParam(0.0 if args.loss == "huber" else 0.5)
        lr = lr.setSolver("l-bfgs").setLoss(args.loss)
        if args.weighted:
            lr = lr.setWeightCol("w")
        bytes_row = 4 + 8 * (d - 1) + 8 + (8 if args.weighted else 0)
        lr.fit(df)  # warm-up: allocations, first-touch
        ctx.enable_pass_timing(True)
        model = lr.fit(df)
        times = ctx.pass_times()
        ctx.enable_pass_timing(False)
        # the evaluation is not an aggregation pass of the fit: wall clock around the call (ms-long passes); the summary
        # is lazy, reading a metric runs the pass
        model.evaluate(df).rootMeanSquaredError()
        t0 = time.perf_counter()
        for _ in range(3):
            model.evaluate(df).rootMeanSquaredError()
        ev = [(time.perf_counter() - t0) * 1e3 / 3]
        big = [t for t in times if t > 0.05 * max(times)]  # drops the 1-CTA column-heads kernel
        summ, lsq = big[0], big[1:]
        frac = lambda ms: bytes_row * rows / (ms * 1e6) / peak  # noqa: E731
        print(json.dumps({"d": d, "rows": rows, "weighted": args.weighted, "loss": args.loss, "bytes_per_row": bytes_row,
                          "summarizer_ms": summ, "summarizer_frac": frac(summ),
                          "lsq_ms_avg": float(np.mean(lsq)) if lsq else None, "lsq_frac": frac(float(np.mean(lsq))) if lsq else None,
                          "lsq_passes": len(lsq), "evaluate_ms": sum(ev), "evaluate_kernels": len(ev), "evaluate_frac": frac(sum(ev)),
                          "all_pass_ms": times}), flush=True)
        del df, model, cols, g, p, x


if __name__ == "__main__":
    main()
