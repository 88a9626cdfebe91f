import base64
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def load_golden(name):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden():
    return load_golden


@pytest.fixture(scope="session")
def lab_csv():
    def get(name):
        return base64.b64decode(load_golden("dataset_" + name)["csv_b64"])
    return get


@pytest.fixture(scope="session")
def pkg():
    import sparkdq4ml_b200
    return sparkdq4ml_b200


@pytest.fixture(scope="session")
def ctx(pkg):
    """One dq4ml context on cuda:0 for the GPU tests.  No fallback: a missing GPU or library is a failure."""
    c = pkg.Context(0)
    yield c
    c.close()


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(1e-300, np.maximum(np.abs(a), np.abs(b))))) if a.size else 0.0


def scale_err(a, b):
    """max |a - b| relative to the largest magnitude in the reference vector: the measure for coefficient vectors, whose
    small entries (an L1 term drives some to, or next to, zero) have no meaningful element-wise relative error."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    if not a.size:
        return 0.0
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-300))


@pytest.fixture(scope="session")
def parity_log():
    """Appends measured parity errors to gpurun_out/parity_errors.jsonl (scratch; summarised under profiles/)."""
    path = os.path.join(ROOT, "gpurun_out", "parity_errors.jsonl")

    def log(test, what, errs):
        try:
            os.makedirs(os.path.dirname(path), exist_ok=True)
            with open(path, "a") as f:
                f.write(json.dumps({"test": test, "what": what, "errors": [float(e) for e in np.atleast_1d(errs)]}) + "\n")
        except OSError:
            pass
    return log


# Tolerance of north_star for fp64 kernels: 1e-9 relative on coefficients / intercept / objective / predictions.
RTOL = 1e-9
