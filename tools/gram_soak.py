"""Soak of the fp64 Gram kernel: repeated passes over one large table must return bit-identical moments (the row
slices are summed in a fixed order whatever the timing of the producer / MMA warps), plain and weighted, with a
ragged row count.    python tools/gram_soak.py [rows] [reps]"""
import os, sys, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import sparkdq4ml_b200 as pkg
from sparkdq4ml_b200 import ml

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_019
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 12
ctx = pkg.Context(0)
out = {}
for d in (256, 130, 65):
    g, p, x = ctx.synth_generate(rows, d=d)
    feats = [g] + [x[k] for k in range(d - 1)]
    first = None
    same = 0
    for r in range(reps):
        got = ml.wls_moments(ctx, feats, p, dq=(g, p))
        if first is None:
            first = got
        same += int(got.tobytes() == first.tobytes())
    out[f"d{d}_plain_identical"] = f"{same}/{reps}"
    assert same == reps, (d, same)
    assert np.isfinite(first).all() and first[0] > 0.9 * rows
    del g, p, x, feats
    torch.cuda.empty_cache()
print(json.dumps({"rows": rows, "reps": reps, **out}))
