#!/usr/bin/env python
"""bench.py -- rows/sec of DQ + LinearRegression fit (BASELINE.json metric) on N B200s.

Workload (BASELINE.json configs[1]): synthetic `guest:int32, price:f64` rows (100 M per GPU, weak scaling), the two DQ
rules + filters fused into the single WeightedLeastSquares aggregation pass, single-feature lasso
(maxIter 40, regParam 1, elasticNetParam 1 -- DataQuality4MachineLearningApp.java:120-123), OWL-QN on the host.
One step = one complete `LinearRegression.fit` through the C ABI.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--rows R]            (torchrun launches N > 1)
  python bench.py --impl reference ...    times the CPU restatement of the reference (oracle/) on the host cores
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "rows/sec DQ+LinearRegression fit"
UNIT = "rows/s"
LAB = dict(max_iter=40, reg_param=1.0, elastic_net_param=1.0)
BYTES_PER_ROW = 12  # int32 guest + f64 price, read once by the fused pass
FALLBACK_HBM_GBS = 6650.0  # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent


def config(rows, extra=None):
    c = {"workload": "configs[1]: synthetic guest/price rows, DQ rules fused + single-feature lasso LinearRegression "
                     "(maxIter=40 regParam=1 elasticNet=1, WLS + OWL-QN)",
         "rows_per_gpu": rows, "features": 1, "l2_policy": "inputs (12 B/row x rows_per_gpu) larger than the 126 MB L2",
         "anomaly_rates": {"p_low": 0.01, "p_high": 0.01}, "seeds": {"data": 20200601, "anomalies": 7}}
    if extra:
        c.update(extra)
    return c


# ------------------------------------------------------------------------------------------------ CPU legs (oracle)

def cpu_fit_rate(sample_rows, min_seconds, max_reps, threads=None, warmup=0):
    """Times oracle fits (fused DQ + WLS moments pass with OpenMP, then the Breeze-style host solve).  Returns
    (rows/s, threads, reps, seconds, list of per-rep seconds)."""
    from oracle import passes, spark_ml

    threads = threads or passes.max_threads()
    guest, price, _ = passes.synth_generate(sample_rows, d=1, threads=threads)
    params = spark_ml.LRParams(**LAB)
    for _ in range(warmup):
        spark_ml.linear_regression_fit([guest], price, params, dq=(guest, price), threads=threads)
    times = []
    t_all = time.perf_counter()
    while len(times) < max_reps:
        t0 = time.perf_counter()
        m = spark_ml.linear_regression_fit([guest], price, params, dq=(guest, price), threads=threads)
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_all >= min_seconds:
            break
    total = sum(times)
    assert m.total_iterations >= 1
    return sample_rows * len(times) / total, threads, len(times), total, times


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample = min(args.rows, args.cpu_sample_rows)
    from oracle import passes, spark_ml

    threads = passes.max_threads()
    guest, price, _ = passes.synth_generate(sample, d=1, threads=threads)
    params = spark_ml.LRParams(**LAB)

    def step():
        return spark_ml.linear_regression_fit([guest], price, params, dq=(guest, price), threads=threads)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = sample * args.steps / dt
    what = f"{sample} rows per step of the same generator/seeds (bounded sample of the {args.rows}-row workload)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config(args.rows, {"cpu_sample_rows": sample}),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": what,
                         "note": "Spark local[*] cannot run here (no JVM): this is oracle/, the C+Python restatement of "
                                 "the reference's algorithm, OpenMP over row partitions"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU leg

class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "50", "-i", str(index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return None
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, mx, reasons = [], [], set()
        for ln in out.splitlines():
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, flag in zip(self.NAMES, parts[3:7]):
                if flag.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return None
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def measured_hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    except Exception:
        return FALLBACK_HBM_GBS, "B200_PROFILING.md fallback (of fallback)"


def ncu_traffic_per_launch(kernel_key):
    """dram bytes per launch of the dominant kernel from the committed `ncu --set full` summary, if any."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_summary.json")) as f:
            return json.load(f).get(kernel_key, {}).get("dram_bytes_per_launch")
    except Exception:
        return None


def run_ours(args):
    import torch
    import torch.distributed as dist

    import sparkdq4ml_b200 as pkg

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    ctx = pkg.Context(local_rank).bind_torch_distributed()
    n = args.rows
    guest, price, _ = ctx.synth_generate(n, d=1, row0=rank * n)
    df = pkg.Frame(ctx, {"guest": guest, "price": price}).with_fused_dq("guest", "price")
    df = df.withColumn("label", df.col("price"))
    df = pkg.VectorAssembler().setInputCols(["guest"]).setOutputCol("features").transform(df)
    lr = pkg.LinearRegression().setMaxIter(LAB["max_iter"]).setRegParam(LAB["reg_param"]).setElasticNetParam(LAB["elastic_net_param"])
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device(f"cuda:{local_rank}"))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.synchronize()

    def timed(step_fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(steps):
            out = step_fn()
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=f"cuda:{local_rank}")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), out

    # ---- device-resident arm (value): inputs already in HBM
    model = None
    for _ in range(max(args.warmup, 3)):
        model = lr.fit(df)
    # the timed region lasts only milliseconds: start the clock sampler first and keep the same step running untimed
    # for a while so that nvidia-smi (50 ms period) sees the GPU under this load, then go straight into the timed steps
    sampler = ClockSampler(local_rank) if rank == 0 else None
    # every rank must run the same number of steps (each fit contains a collective): rank 0 sizes the pre-roll from
    # its warm-up speed and tells the others
    t0 = time.perf_counter()
    model = lr.fit(df)
    per_step = max(time.perf_counter() - t0, 1e-5)
    pre_steps = torch.tensor([min(int(args.preroll_seconds / per_step), 20000)], dtype=torch.int64, device=f"cuda:{local_rank}")
    if world > 1:
        dist.broadcast(pre_steps, src=0)
    for _ in range(int(pre_steps.item())):
        model = lr.fit(df)
    ctx.enable_pass_timing(True)
    l0 = ctx.launch_count
    ms, model = timed(lambda: lr.fit(df), args.steps)
    launches = ctx.launch_count - l0
    pass_ms, passes = ctx.pass_timing()
    ctx.enable_pass_timing(False)
    clocks = sampler.stop() if sampler is not None else None
    value = world * n * args.steps / (ms / 1e3)

    # ---- end-to-end arm: HOST (pinned) columns through the same C-ABI call; H2D inside the timed region
    gh = guest.cpu().pin_memory()
    ph = price.cpu().pin_memory()
    hdf = pkg.Frame(ctx, {"guest": gh, "price": ph}).with_fused_dq("guest", "price")
    hdf = hdf.withColumn("label", hdf.col("price"))
    hdf = pkg.VectorAssembler().setInputCols(["guest"]).setOutputCol("features").transform(hdf)
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    for _ in range(2):
        hmodel = lr.fit(hdf)
    e2e_ms, hmodel = timed(lambda: lr.fit(hdf), e2e_steps)
    e2e_value = world * n * e2e_steps / (e2e_ms / 1e3)
    assert hmodel.coefficients().tobytes() == model.coefficients().tobytes()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = measured_hbm_peak()
    per_launch_ms = pass_ms / max(passes, 1)
    achieved = BYTES_PER_ROW * n / (per_launch_ms / 1e3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": config(n),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": BYTES_PER_ROW * n, "d2h_bytes_per_step": 6 * 8,
                "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps,
                "note": "pinned host columns -> dq4ml_linreg_fit; H2D staging + pass + D2H of the moments inside the timed region"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "stream_reduce_kernel<WlsOp<1>,1,8,kDqLab> (fused DQ + WLS moments)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic_per_launch("wls_moments_d1_lab"), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": BYTES_PER_ROW * n, "avg_launch_ms": per_launch_ms,
                     "launches_timed": int(passes)},
        "clocks": clocks,
        "fit": {"coefficient": float(model.coefficients()[0]), "intercept": model.intercept(),
                "total_iterations": model.summary().totalIterations(), "rows_after_dq": model.num_rows,
                "solver": model.solver_used},
    }
    if world == 1 and not args.no_cpu_baseline:
        rate, threads, reps, secs, _ = cpu_fit_rate(min(n, args.cpu_sample_rows), args.cpu_seconds, 10**9)
        line["cpu_baseline"] = {
            "value": rate, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{reps} fits of {min(n, args.cpu_sample_rows)} rows (same generator/seeds), {secs:.1f} s of CPU work",
            "note": "oracle/ (C passes with OpenMP + Python Breeze restatement); Spark local[*] cannot run here (no JVM)"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--rows", type=int, default=100_000_000, help="rows per GPU (weak scaling)")
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--cpu-sample-rows", type=int, default=50_000_000)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--preroll-seconds", type=float, default=2.0, help="untimed steps under the clock sampler")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
