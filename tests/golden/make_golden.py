"""Regenerates tests/golden/*.json.  Run in the authoring container only (needs /root/reference):

    python tests/golden/make_golden.py

What is stored, and where each value comes from:
  * dataset_{abstract,small,full}.json -- the reference's three fixtures (data/dataset-*.csv, read by App:52-55),
    carried as base64 so the GPU box (which has no /root/reference) can run config 1; their parsed columns; the
    rows surviving the DQ rules; sha256 of the packed survivors; WLS moments; fits under several hyper-parameter
    sets computed by the oracle (oracle/), and INDEPENDENT cross-checks of the converged optimum: the closed-form
    lasso/OLS solution for d = 1 and a scikit-learn coordinate-descent solve.
  * survey_kats.json -- known-answer values derived independently by the surveyor, typed in by hand.
  * synth_small.json -- d = 4 synthetic case: oracle fits + scikit-learn cross-check.
The reference publishes no golden outputs of its own (no tests): parity is "unpinned" in that sense.
"""
import base64
import hashlib
import json
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))
from oracle import passes, spark_ml  # noqa: E402

REF = "/root/reference/data"

PARAM_SETS = {
    "lab_wls_lasso": dict(max_iter=40, reg_param=1.0, elastic_net_param=1.0),  # App:120-123
    "lab_lbfgs_lasso": dict(max_iter=40, reg_param=1.0, elastic_net_param=1.0, solver="l-bfgs"),
    "ols_cholesky": dict(max_iter=40, reg_param=0.0, elastic_net_param=0.0),
    "ridge_cholesky": dict(max_iter=40, reg_param=0.5, elastic_net_param=0.0),
    "enet_wls": dict(max_iter=40, reg_param=0.7, elastic_net_param=0.5),
    "enet_lbfgs": dict(max_iter=40, reg_param=0.7, elastic_net_param=0.5, solver="l-bfgs"),
    "ols_lbfgs": dict(max_iter=100, reg_param=0.0, elastic_net_param=0.0, solver="l-bfgs"),
    "nostd_enet_wls": dict(max_iter=40, reg_param=0.7, elastic_net_param=0.5, standardization=False),
    "nointercept_lbfgs": dict(max_iter=40, reg_param=0.1, elastic_net_param=0.3, solver="l-bfgs", fit_intercept=False),
}


def sklearn_enet(X, y, reg, alpha, sample_std):
    """Converged optimum of Spark's objective (standardization=true, fitIntercept=true) by coordinate descent."""
    from sklearn.linear_model import ElasticNet, LinearRegression
    n = len(y)
    ddof = 1 if sample_std else 0
    sx = X.std(axis=0, ddof=ddof)
    sy = y.std(ddof=ddof)
    Xs, ys = X / sx, y / sy
    if reg == 0.0:
        m = LinearRegression().fit(Xs, ys)
    else:
        m = ElasticNet(alpha=reg / sy, l1_ratio=alpha, fit_intercept=True, tol=1e-14, max_iter=1000000).fit(Xs, ys)
    coef = m.coef_ * sy / sx
    return coef.tolist(), float(y.mean() - coef @ X.mean(axis=0))


def fit_block(features, label, d):
    out = {}
    for name, kw in PARAM_SETS.items():
        p = spark_ml.LRParams(**kw)
        m = spark_ml.linear_regression_fit(features, label, p)
        pred, met = passes.predict_metrics(features, label, m.coefficients, m.intercept)
        rm = spark_ml.regression_metrics(met, p.fit_intercept)
        out[name] = {
            "params": kw, "coefficients": m.coefficients.tolist(), "intercept": m.intercept,
            "objective_history": [float(v) for v in m.objective_history], "total_iterations": m.total_iterations,
            "solver_used": m.solver_used, "rmse": rm["rmse"], "r2": rm["r2"],
            "predict_40": float(m.coefficients[0] * 40.0 + m.intercept) if d == 1 else None,
        }
        if p.standardization and p.fit_intercept:
            X = np.stack([np.asarray(f, dtype=np.float64) for f in features], axis=1)
            c, b = sklearn_enet(X, np.asarray(label), p.reg_param, p.elastic_net_param, sample_std=(p.solver == "l-bfgs"))
            out[name]["sklearn_coefficients"] = c
            out[name]["sklearn_intercept"] = b
    return out


def main():
    for name in ("abstract", "small", "full"):
        raw = open(os.path.join(REF, f"dataset-{name}.csv"), "rb").read()
        g, p = passes.csv_parse(raw)
        # independent parse with Python's own float()
        recs = [r for r in raw.decode().replace("\r\n", "\n").replace("\r", "\n").split("\n") if r]
        assert [int(r.split(",")[0]) for r in recs] == g.tolist()
        assert [float(r.split(",")[1]) for r in recs] == p.tolist()
        go, po, n1 = passes.dq_filter(g, p)
        packed = np.zeros(len(go), dtype=[("g", "<i4"), ("p", "<f8")])
        packed["g"], packed["p"] = go, po
        x, y = go.astype(np.float64), po
        n = len(y)
        # closed-form lasso in scaled space (population sigma): beta* = sign(rho) max(|rho| - lambda/sigma_y, 0)
        sx, sy = x.std(), y.std()
        rho = float(((x - x.mean()) * (y - y.mean())).mean() / (sx * sy))
        c = math.copysign(max(abs(rho) - 1.0 / sy, 0.0), rho)
        coef_cf = c * sy / sx
        doc = {
            "source": f"data/dataset-{name}.csv of the reference (read at App:52-55)",
            "csv_b64": base64.b64encode(raw).decode(), "guest": g.tolist(), "price": p.tolist(),
            "n": int(len(g)), "n_after_rule1": int(n1), "n_after_rule2": int(len(go)),
            "survivor_guest": go.tolist(), "survivor_price": po.tolist(),
            "survivor_sha256": hashlib.sha256(packed.tobytes()).hexdigest(),
            "wls_moments": passes.wls_moments([go], po).tolist(),
            "closed_form_lasso": {"coefficient": coef_cf, "intercept": float(y.mean() - coef_cf * x.mean())},
            "fits": fit_block([go], po, 1),
        }
        json.dump(doc, open(os.path.join(HERE, f"dataset_{name}.json"), "w"), indent=1)
        print(name, n, doc["survivor_sha256"][:8], doc["fits"]["lab_wls_lasso"]["coefficients"])

    guest, price, extra = passes.synth_generate(4000, d=4, p_low=0.02, p_high=0.02)
    go_mask = np.array([passes.lib().orc_dq_keep(int(a), float(b)) for a, b in zip(guest, price)], dtype=bool)
    feats = [guest[go_mask]] + [np.ascontiguousarray(extra[k][go_mask]) for k in range(3)]
    lab = price[go_mask]
    doc = {"n": 4000, "d": 4, "p_low": 0.02, "p_high": 0.02, "seed_data": 20200601, "seed_anom": 7,
           "n_survivors": int(go_mask.sum()),
           "head_guest": guest[:16].tolist(), "head_price": price[:16].tolist(), "head_extra": extra[:, :16].tolist(),
           "digest": [str(v) for v in passes.rows_digest(guest, price)],
           "fits": fit_block(feats, lab, 4)}
    json.dump(doc, open(os.path.join(HERE, "synth_small.json"), "w"), indent=1)
    print("synth", doc["n_survivors"], doc["fits"]["enet_lbfgs"]["coefficients"], doc["fits"]["enet_lbfgs"]["sklearn_coefficients"])


if __name__ == "__main__":
    main()
