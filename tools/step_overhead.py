#!/usr/bin/env python
"""Where the time of one bench step (a whole LinearRegression.fit on config 2) goes besides the kernel: per-call cost of
the ctypes boundary, of a fit over a table so small that the kernel is negligible (host path + launch + hand-over), and
the wall / event time of the full-size fit.  Prints one JSON line."""
import json
import os
import sys
import time

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)


def main():
    import sparkdq4ml_b200 as pkg
    ctx = pkg.Context(0)
    L = pkg.capi.lib()
    out = {"epilogue_env": os.environ.get("DQ4ML_EPILOGUE", "on")}
    n_calls = 200000
    t0 = time.perf_counter()
    for _ in range(n_calls):
        L.dq4ml_abi_version()
    out["ctypes_trivial_call_us"] = (time.perf_counter() - t0) / n_calls * 1e6
    lr = pkg.LinearRegression().setMaxIter(40).setRegParam(1.0).setElasticNetParam(1.0)
    for n in (1024, 100_000_000):
        g, p, _ = ctx.synth_generate(n, d=1)
        df = pkg.Frame(ctx, {"guest": g, "price": p}).with_fused_dq("guest", "price")
        df = df.withColumn("label", df.col("price"))
        df = pkg.VectorAssembler().setInputCols(["guest"]).setOutputCol("features").transform(df)
        reps = 3000 if n < 10**6 else 300
        for _ in range(20):
            lr.fit(df)
        ctx.enable_pass_timing(True)
        t0 = time.perf_counter()
        for _ in range(reps):
            lr.fit(df)
        wall = (time.perf_counter() - t0) / reps * 1e6
        ms, passes = ctx.pass_timing()
        ctx.enable_pass_timing(False)
        t0 = time.perf_counter()
        for _ in range(reps):
            lr.fit(df)
        wall_untimed = (time.perf_counter() - t0) / reps * 1e6
        out[f"fit_n{n}"] = {"wall_us_per_fit_with_pass_events": wall, "wall_us_per_fit": wall_untimed,
                            "kernel_event_us": ms / max(passes, 1) * 1e3}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
