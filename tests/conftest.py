import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    """`gpu`-marked tests need a CUDA device and the built library: without them they are skipped (a plain `pytest tests`
    on a CPU-only host stays green), never run against some fallback -- there is none."""
    try:
        import torch
        have_gpu = torch.cuda.is_available()
    except Exception:
        have_gpu = False
    have_lib = os.path.exists(os.path.join(ROOT, "pythonphot_b200", "csrc", "libbphot.so"))
    if have_gpu and have_lib:
        return
    why = "no CUDA device" if not have_gpu else "pythonphot_b200/csrc/libbphot.so is not built"
    skip = pytest.mark.skip(reason="gpu test: " + why)
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


def assert_close(a, b, rtol=1e-9, atol=0.0, what=""):
    """NaN pattern must match exactly; finite values within rtol (relative) + atol."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, "%s: shape %s vs %s" % (what, a.shape, b.shape)
    na, nb = np.isnan(a), np.isnan(b)
    assert np.array_equal(na, nb), "%s: NaN pattern differs at %s" % (what, np.argwhere(na != nb)[:5].tolist())
    ia, ib = np.isinf(a), np.isinf(b)
    assert np.array_equal(ia, ib) and np.array_equal(a[ia], b[ib]), "%s: inf pattern differs" % what
    ok = ~(na | ia)
    err = np.abs(a[ok] - b[ok])
    tol = atol + rtol * np.abs(b[ok])
    bad = err > tol
    if bad.any():
        k = np.argmax(err - tol)
        raise AssertionError("%s: %d/%d outside tol; worst got %r want %r" % (what, bad.sum(), ok.sum(), a[ok][k], b[ok][k]))
