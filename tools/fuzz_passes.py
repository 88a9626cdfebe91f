"""Randomised parity sweep of the streaming passes against the oracle: loss/gradient (every width 1..32, fused DQ in the
lab layout, separate, or off; host or device columns) -- exercises every dispatch branch of stream_launch.cu.
    python tools/fuzz_passes.py [cases] [seed]"""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import sparkdq4ml_b200 as pkg
from sparkdq4ml_b200 import ml
from oracle import passes

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 60
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 7)
ctx = pkg.Context(0)
worst = 0.0
for c in range(cases):
    d = int(rng.integers(1, 33))
    n = int(rng.choice([1, 7, 255, 1024, 6144, 6145, 40_000, 100_003, 250_001]))
    mode = int(rng.integers(0, 3))  # 0 = no DQ, 1 = lab layout, 2 = separate copies of the DQ columns
    guest, price, extra = passes.synth_generate(n, d=d, p_low=0.04, p_high=0.04, seed_data=int(rng.integers(1, 1 << 30)))
    feats = [guest] + [np.ascontiguousarray(extra[k]) for k in range(d - 1)]
    if rng.random() < 0.3 and d > 1:  # all-fp64 table: move guest out of the features
        feats = [np.ascontiguousarray(extra[0] * 3.0 + 1.0)] + feats[1:]
    dev = [torch.from_numpy(np.ascontiguousarray(f)).cuda() for f in feats]
    dprice, dguest = torch.from_numpy(price).cuda(), torch.from_numpy(guest).cuda()
    coef = rng.normal(size=d)
    fstd, fmean = rng.uniform(0.5, 2.0, size=d), rng.normal(size=d)
    dq_host = None if mode == 0 else (guest, price)
    if mode == 0:
        dq_dev = None
    elif mode == 1 and feats[0] is guest:
        dq_dev = (dev[0], dprice)
    else:
        dq_dev = (dguest.clone(), dprice.clone())
    rl, rw, rg = passes.lsq_loss_grad(feats, price, coef, fstd, fmean, 20.0, 100.0, True, dq=dq_host, threads=4)
    gl, gw, gg = ml.lsq_loss_grad(ctx, dev, dprice, coef, fstd, fmean, 20.0, 100.0, True, dq=dq_dev)
    assert gw == rw, (c, d, n, mode, gw, rw)
    e = max(abs(gl - rl) / max(abs(rl), 1e-300), float(np.max(np.abs(gg - rg)) / max(np.abs(rg).max(), 1e-300)))
    worst = max(worst, e)
    assert e < 1e-9, (c, d, n, mode, e)
print(json.dumps({"cases": cases, "worst_relative_error": worst}))
