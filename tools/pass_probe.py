"""Times one loss/gradient pass (dq4ml_lsq_loss_grad) over a synthetic table: python tools/pass_probe.py d rows"""
import os, sys, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparkdq4ml_b200 as pkg
from sparkdq4ml_b200 import ml

d, rows = int(sys.argv[1]), int(sys.argv[2])
ctx = pkg.Context(0)
g, p, x = ctx.synth_generate(rows, d=d)
feats = [g] + [x[k] for k in range(d - 1)]
coef, fstd, fmean = np.full(d, 0.1), np.full(d, 1.5), np.full(d, 0.7)
for rep in range(3):
    ctx.enable_pass_timing(True)
    out = ml.lsq_loss_grad(ctx, feats, p, coef, fstd, fmean, 20.0, 100.0, True, dq=(g, p))
    ms, cnt = ctx.pass_timing()
    ctx.enable_pass_timing(False)
bytes_row = 4 + 8 * d
print(json.dumps({"d": d, "rows": rows, "kernels": cnt, "total_ms": ms, "pass_ms": ms / cnt, "gbs": bytes_row * rows / (ms / cnt) / 1e6, "count": float(out[1])}))
