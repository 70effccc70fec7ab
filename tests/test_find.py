"""find.find -- oracle and CUDA path against the reference's golden vectors."""
import contextlib
import io

import numpy as np
import pytest

from tests.conftest import assert_close, load_golden

CASES = {  # name -> hmin, fwhm, roundlim, sharplim (tests/golden/gen_golden_psf_find.py)
    "fwhm4": (60.0, 4.1, [-1.0, 1.0], [0.2, 1.0]),
    "fwhm2p5_loose": (40.0, 2.5, [-2.0, 2.0], [0.1, 1.2]),
    "fwhm9_bigbox": (80.0, 9.0, [-1.0, 1.0], [0.2, 1.0]),
    "none_found": (1e9, 4.1, [-1.0, 1.0], [0.2, 1.0]),
    "with_nans": (60.0, 4.1, [-1.0, 1.0], [0.2, 1.0]),
}
NAMES = ("x", "y", "flux", "sharp", "round")


def _image(g, name):
    img = g["image"]
    if name == "with_nans":
        img = img.copy()
        img.reshape(-1)[g["nan_idx"]] = np.nan
    return img


def _compare(res, g, name, rtol):
    want = [g["%s__%s" % (name, nm)] for nm in NAMES]
    if want[0].size == 0:
        assert res is None
        return
    assert res is not None and len(res) == 5
    assert res[0].shape == want[0].shape, "%s: %d stars found, reference found %d" % (name, res[0].size, want[0].size)
    for nm, a, w in zip(NAMES, res, want):
        # positions are pixel index + offset: compare to 1e-9 px absolute on top of the relative bound
        assert_close(a, w, rtol=rtol, atol=1e-9 if nm in ("x", "y", "round") else 0.0, what="%s %s" % (name, nm))


@pytest.mark.parametrize("name", sorted(CASES))
def test_find_oracle_matches_reference_golden(name):
    from oracle import find_oracle
    g = load_golden("golden_find.npz")
    hmin, fwhm, rl, sl = CASES[name]
    _compare(find_oracle.find(_image(g, name), hmin, fwhm, rl, sl), g, name, rtol=0.0 if name != "none_found" else 0.0)


def test_find_oracle_counts_add_up():
    from oracle import find_oracle
    g = load_golden("golden_find.npz")
    res, cnt = find_oracle.find(g["image"], 60.0, 4.1, return_counts=True)
    assert cnt["ngood"] == res[0].size + cnt["badsharp"] + cnt["badround"] + cnt["badcntrd"]


# ----------------------------------------------------------------------------- CUDA path
def _quiet_find(*a, **k):
    from pythonphot_b200 import find
    with contextlib.redirect_stdout(io.StringIO()):
        return find.find(*a, verbose=False, **k)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_find_gpu_matches_reference_golden(name):
    g = load_golden("golden_find.npz")
    hmin, fwhm, rl, sl = CASES[name]
    _compare(_quiet_find(_image(g, name), hmin, fwhm, rl, sl), g, name, rtol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_find_gpu_matches_oracle_on_large_field(dtype):
    """2048 x 1489 SDSS-like frame (BASELINE.json configs[0] stand-in), ~4k stars, fp32 and fp64 pixels."""
    import torch
    from oracle import find_oracle
    from pythonphot_b200 import synth
    from pythonphot_b200.find import find_device
    f = synth.make_field(1489, 2048, 4000, seed=1001, dtype=dtype, margin=2.0, edge_frac=0.02, nan_frac=0.0, with_noise=False)
    want, wc = find_oracle.find(f["image"].astype(np.float64), 50.0, 4.1, return_counts=True)
    cols, cnt = find_device(torch.from_numpy(f["image"]).cuda(), 50.0, 4.1, max_out=64)  # small on purpose: exercises the retry
    got = cols.cpu().numpy()
    assert cnt["ngood"] == wc["ngood"] and cnt["nstars"] == want[0].size
    assert (cnt["badsharp"], cnt["badround"], cnt["badcntrd"]) == (wc["badsharp"], wc["badround"], wc["badcntrd"])
    for i, nm in enumerate(NAMES):
        assert_close(got[i], want[i], rtol=1e-9, atol=1e-9 if nm in ("x", "y", "round") else 0.0, what=nm)
    # stars come out in row-major order of their peak pixel (np.where order, find.py:211)
    peak = np.rint(got[1] - 0).astype(int)
    assert want[0].size > 2000


@pytest.mark.gpu
def test_find_gpu_host_entry_and_errors():
    import ctypes as C
    from pythonphot_b200 import _lib
    g = load_golden("golden_find.npz")
    img = np.ascontiguousarray(g["image"])
    n = g["fwhm4__x"].size
    cols = [np.empty(n) for _ in range(5)]
    counts = (C.c_int * 5)()
    rc = _lib.lib().bphot_find_host(_lib.nptr(img), _lib.F64, img.shape[0], img.shape[1], 60.0, 4.1, -1.0, 1.0, 0.2, 1.0, n,
                                    *[_lib.nptr(c) for c in cols], C.cast(counts, C.c_void_p), 0)
    assert rc == 0 and counts[0] == n
    for nm, c in zip(NAMES, cols):
        assert_close(c, g["fwhm4__" + nm], rtol=1e-9, atol=1e-9, what=nm)
    rc = _lib.lib().bphot_find_host(_lib.nptr(img), _lib.F64, img.shape[0], img.shape[1], 60.0, 4.1, -1.0, 1.0, 0.2, 1.0, n - 1,
                                    *[_lib.nptr(c) for c in cols], C.cast(counts, C.c_void_p), 0)
    assert rc == _lib.ERR_CAPACITY and counts[0] == n
    rc = _lib.lib().bphot_find_host(_lib.nptr(img), _lib.F64, img.shape[0], img.shape[1], 60.0, 0.3, -1.0, 1.0, 0.2, 1.0, n,
                                    *[_lib.nptr(c) for c in cols], C.cast(counts, C.c_void_p), 0)
    assert rc == -1 and b"FWHM" in _lib.lib().bphot_last_error()
