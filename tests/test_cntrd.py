"""cntrd.cntrd -- oracle and CUDA path against golden vectors from the unmodified reference."""
import contextlib
import io

import numpy as np
import pytest

from tests.conftest import assert_close, load_golden

CASES = [("fwhm4", 4.1, 0, False), ("fwhm4_extend3", 4.1, 3, False), ("fwhm4_extend_xy", 4.1, (1, 4), False),
         ("fwhm7_keepcenter", 7.0, 0, True), ("fwhm2", 2.0, 0, False)]
FRAMES = ["", "rounded__"]


def _frame(g, tag):
    return g["image"] if tag == "" else np.round(g["image"] / 50.0) * 50.0


@pytest.mark.parametrize("tag", FRAMES)
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_cntrd_oracle_matches_reference_golden(tag, case):
    from oracle import cntrd_oracle
    g = load_golden("golden_cntrd.npz")
    name, fwhm, ext, keep = case
    with np.errstate(all="ignore"):
        xc, yc, st = cntrd_oracle.cntrd(_frame(g, tag), g["x"], g["y"], fwhm, ext, keep)
    assert np.array_equal(xc, g[tag + name + "__xcen"], equal_nan=True)
    assert np.array_equal(yc, g[tag + name + "__ycen"], equal_nan=True)
    assert np.array_equal(st != 0, xc == -1)


@pytest.mark.gpu
@pytest.mark.parametrize("tag", FRAMES)
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_cntrd_gpu_matches_reference_golden(tag, case):
    from pythonphot_b200 import cntrd
    g = load_golden("golden_cntrd.npz")
    name, fwhm, ext, keep = case
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        xc, yc = cntrd.cntrd(_frame(g, tag), g["x"], g["y"], fwhm, verbose=True, extendbox=ext, keepcenter=keep)
    wx, wy = g[tag + name + "__xcen"], g[tag + name + "__ycen"]
    assert np.array_equal(xc == -1, wx == -1), "different stars rejected"
    assert_close(xc, wx, rtol=1e-9, atol=1e-9, what=name + " xcen")
    assert_close(yc, wy, rtol=1e-9, atol=1e-9, what=name + " ycen")
    assert len(buf.getvalue().splitlines()) == int((wx == -1).sum())  # one message per rejected position, as the reference prints


@pytest.mark.gpu
def test_cntrd_gpu_scalar_and_fp32_and_oracle_statuses():
    import torch
    from oracle import cntrd_oracle
    from pythonphot_b200 import cntrd, synth
    g = load_golden("golden_cntrd.npz")
    k = int(g["scalar__k"])
    xs, ys = cntrd.cntrd(g["image"], float(g["x"][k]), float(g["y"][k]), 4.1, verbose=False)
    assert isinstance(xs, float) and abs(xs - g["scalar__xy"][0]) < 1e-9 and abs(ys - g["scalar__xy"][1]) < 1e-9
    # a 20k-star fp32 field against the oracle (widened pixels), status codes included
    f = synth.make_field(1024, 1024, 20000, seed=43, dtype=np.float32, margin=2.0, edge_frac=0.03, nan_frac=1e-4, with_noise=False)
    rng = np.random.default_rng(44)
    x, y = f["x"] + rng.normal(0, 0.7, f["x"].size), f["y"] + rng.normal(0, 0.7, f["y"].size)
    with np.errstate(all="ignore"):
        wx, wy, wst = cntrd_oracle.cntrd(f["image"].astype(np.float64), x, y, 4.1, 1, False)
    xc, yc, st = cntrd.cntrd_device(torch.from_numpy(f["image"]).cuda(), torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), 4.1, 1, False)
    assert np.array_equal(st.cpu().numpy(), wst)
    assert_close(xc.cpu().numpy(), wx, rtol=1e-9, atol=1e-9, what="xcen")
    assert_close(yc.cpu().numpy(), wy, rtol=1e-9, atol=1e-9, what="ycen")
    assert (wst == 0).sum() > 10000
