import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    """-> (arrays dict, meta dict) of one fixture written by oracle/make_golden.py."""
    with np.load(os.path.join(GOLDEN_DIR, name + ".npz")) as z:
        arrays = {k: z[k] for k in z.files if k != "__meta__"}
        meta = json.loads(str(z["__meta__"]))
    return arrays, meta


def golden_names(prefix):
    return sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.startswith(prefix) and f.endswith(".npz"))


def check_against(arrays, key, value, rtol, atol):
    """Compare ``value`` with fixture entry ``key`` (whole, or its strided sub-sample + sum)."""
    value = np.asarray(value, dtype=np.float64)
    if key in arrays:
        np.testing.assert_allclose(value, arrays[key], rtol=rtol, atol=atol, err_msg=key)
        return
    sub = arrays[key + "__sub"]
    stride = int(round(value.size / sub.size)) if sub.size else 1
    got = value.reshape(-1)[::stride]
    if got.size != sub.size:       # stride was recorded implicitly; recover it exactly
        for s in range(1, 4096):
            if value.reshape(-1)[::s].size == sub.size:
                got = value.reshape(-1)[::s]
                break
    np.testing.assert_allclose(got, sub, rtol=rtol, atol=atol, err_msg=key + "__sub")
    np.testing.assert_allclose(value.sum(), arrays[key + "__sum"], rtol=max(rtol, 1e-9),
                               atol=atol * value.size, err_msg=key + "__sum")


def check_means_within_standard_error(got, want, what, sigmas=3.0):
    """Two samples of one statistic (e.g. 64 dropout seeds each): |mean difference| <= sigmas * combined
    standard error, and the spreads within a factor 1.6 of each other."""
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    se = np.sqrt(got.var(ddof=1) / got.size + want.var(ddof=1) / want.size)
    assert abs(got.mean() - want.mean()) <= sigmas * se, (what, got.mean(), want.mean(), se)
    ratio = got.std(ddof=1) / want.std(ddof=1)
    assert 1 / 1.6 < ratio < 1.6, (what, "spread", got.std(ddof=1), want.std(ddof=1))


@pytest.fixture
def workdir(tmp_path, monkeypatch):
    """The model class persists under ./models relative to cwd (as the reference does)."""
    monkeypatch.chdir(tmp_path)
    return tmp_path
