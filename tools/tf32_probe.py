"""Development probe for the tcgen05 Gram tier: moments of the tf32x3 tier against the fp64 tier on the same device
columns, error measured relative to sqrt(G_ii G_jj).  Usage: python tools/tf32_probe.py [d n]..."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparkdq4ml_b200 as pkg  # noqa: E402
from sparkdq4ml_b200 import ml  # noqa: E402


def unpack(m, d):
    tri = m[5 + 2 * d:]
    G = np.zeros((d, d))
    q = 0
    for j in range(d):
        for i in range(j + 1):
            G[i, j] = G[j, i] = tri[q]
            q += 1
    return G


def main():
    shapes = [(int(a), int(b)) for a, b in zip(sys.argv[1::2], sys.argv[2::2])] or [(9, 16), (16, 4096), (128, 5000),
                                                                                   (130, 5003), (256, 20011)]
    ctx = pkg.Context(0)
    for d, n in shapes:
        rng = np.random.default_rng(d * 1000 + n)
        cols = [torch.from_numpy(rng.integers(1, 36, n).astype(np.int32)).cuda()]
        cols += [torch.from_numpy(rng.uniform(0, 1 + k, n)).cuda() for k in range(1, d)]
        y = torch.from_numpy(rng.normal(50, 20, n)).cuda()
        ctx.set_gram_tier("fp64")
        ref = ml.wls_moments(ctx, cols, y)
        ctx.set_gram_tier("tf32x3")
        t0 = time.time()
        got = ml.wls_moments(ctx, cols, y)
        dt = time.time() - t0
        ctx.set_gram_tier("fp64")
        Gr, Gg = unpack(ref, d), unpack(got, d)
        scale = np.sqrt(np.outer(np.diag(Gr), np.diag(Gr)))
        err = np.abs(Gg - Gr) / scale
        mean = ref[5:5 + d] / ref[1]
        cov_r = Gr / ref[1] - np.outer(mean, mean)
        cov_g = Gg / got[1] - np.outer(got[5:5 + d] / got[1], got[5:5 + d] / got[1])
        sd = np.sqrt(np.diag(cov_r))
        cerr = np.abs(cov_g - cov_r) / np.outer(sd, sd)
        head = np.abs(got[:5 + 2 * d] - ref[:5 + 2 * d]) / np.maximum(np.abs(ref[:5 + 2 * d]), 1e-300)
        print(f"d={d} n={n}: raw-moment err max {err.max():.3e} at {np.unravel_index(err.argmax(), err.shape)}; "
              f"covariance err max {cerr.max():.3e}; scalar/border err max {head.max():.3e}; {dt * 1e3:.1f} ms", flush=True)
        if err.max() > 1e-3:
            np.set_printoptions(precision=4, linewidth=200)
            print("ref[:6,:6]\n", Gr[:6, :6], "\ngot[:6,:6]\n", Gg[:6, :6])
            if d > 128:
                print("ref[126:132,126:132]\n", Gr[126:132, 126:132], "\ngot\n", Gg[126:132, 126:132])


if __name__ == "__main__":
    main()
