"""getpsf.getpsf against golden vectors from the unmodified reference.

CPU leg: the host orchestration of pythonphot_b200/getpsf.py with the three CUDA primitives it calls
(daoerf, rinter, pkfit) replaced by the oracle's CPU restatements -- checks the control flow, the 5x5
solve, the NaN patching and the FITS file without a GPU.  GPU leg: the real thing."""
import os

import numpy as np
import pytest

from tests.conftest import assert_close, load_golden

CASES = ["r8_f3", "r6_f2p5", "single_star", "edge_star_first"]
PARAMS = {"r8_f3": (8, 3), "r6_f2p5": (6, 2.5), "single_star": (8, 3), "edge_star_first": (8, 3), "fitrad_too_big": (4, 5)}


def _run(getpsf_mod, g, name, psfname=None):
    psfrad, fitrad = PARAMS[name]
    import contextlib
    import io
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        out = getpsf_mod.getpsf(g["image"], g["x"], g["y"], g["mag"], g["sky"], float(g["ronois"]), float(g["phpadu"]),
                                g[name + "__idpsf"], psfrad, fitrad, psfname, verbose=False)
    return out, buf.getvalue()


def _check(out, g, name):
    gauss, psf, psfmag = out
    assert_close(gauss, g[name + "__gauss"], rtol=1e-9, what=name + " gauss")
    assert_close(psfmag, g[name + "__psfmag"], rtol=1e-9, what=name + " psfmag")
    want = g[name + "__psf"]
    assert_close(psf, want, rtol=1e-9, atol=1e-9 * np.abs(want).max(), what=name + " psf table")


class _OraclePrimitives(object):
    """Stand-ins with the signatures of pythonphot_b200.daoerf / rinter / pkfit, computed by the CPU oracle."""

    def __init__(self):
        from oracle import oracle
        self.o = oracle
        oracle.build()

    def daoerf(self, x, y, a, device=None):
        return self.o.daoerf(x, y, a)

    def rinter(self, p, x, y, deriv=True, device=None):
        z, dx, dy = self.o.rinter(p, x, y)
        return (z, dx, dy) if deriv else z

    def pkfit_class(self, f, gauss, psf, ronois, phpadu, device=None):
        o = self.o

        class _PK(object):
            def pkfit(self, scale, x, y, sky, radius, debug=False, xyout=False):
                r = o.pkfit(o.PKFIT, f, gauss, psf, ronois, phpadu, None, None, scale, x, y, sky, radius)
                return tuple((int(r[k][0]) if k == "niter" else float(r[k][0])) for k in ("errmag", "chi", "sharp", "niter", "scale", "x", "y"))
        return _PK()


@pytest.fixture
def getpsf_with_oracle(monkeypatch):
    from pythonphot_b200 import getpsf as gp
    prim = _OraclePrimitives()
    ns = type("ns", (), {})
    d, r, p = ns(), ns(), ns()
    d.daoerf, r.rinter, p.pkfit_class = prim.daoerf, prim.rinter, prim.pkfit_class
    monkeypatch.setattr(gp, "daoerf", d)
    monkeypatch.setattr(gp, "rinter", r)
    monkeypatch.setattr(gp, "pkfit", p)
    return gp


@pytest.mark.parametrize("name", CASES)
def test_getpsf_host_logic_matches_reference_golden(getpsf_with_oracle, name):
    g = load_golden("golden_getpsf.npz")
    out, log = _run(getpsf_with_oracle, g, name)
    _check(out, g, name)
    assert log.count("too near edge") + log.count("No convergence") == int(g[name + "__nskipped"])


def test_getpsf_failure_tuple_and_fits_roundtrip(getpsf_with_oracle, tmp_path):
    g = load_golden("golden_getpsf.npz")
    (gauss, psf, psfmag), _ = _run(getpsf_with_oracle, g, "fitrad_too_big")
    assert psf == -1 and psfmag == -1 and np.array_equal(gauss, g["fitrad_too_big__gauss"])
    path = os.path.join(str(tmp_path), "psf.fits")
    (gauss, psf, psfmag), _ = _run(getpsf_with_oracle, g, "single_star", psfname=path)
    from pythonphot_b200 import _fits
    data, hdr = _fits.read_image(path)
    assert np.array_equal(data, psf) and data.dtype == np.float64
    assert [hdr["GAUSS%d" % (i + 1)] for i in range(5)] == [float(v) for v in gauss]
    assert hdr["PSFMAG"] == float(psfmag) and hdr["PSFRAD"] == 8.0 and hdr["FITRAD"] == 3.0 and hdr["NSTARS"] == 1
    assert os.path.getsize(path) % 2880 == 0


def test_fits_roundtrip_dtypes(tmp_path):
    from pythonphot_b200 import _fits
    rng = np.random.default_rng(3)
    for arr in (rng.normal(size=(5, 7)), rng.normal(size=(4, 3)).astype(np.float32), rng.integers(0, 255, (6, 2)).astype(np.uint8),
                rng.integers(-9, 9, (3, 3)).astype(np.int32)):
        p = os.path.join(str(tmp_path), "a.fits")
        _fits.write_image(p, arr, [("OBJECT", "it's a test"), ("EXPTIME", 53.9), ("FLAG", True), ("N", 7), ("BAD", float("nan"))])
        back, hdr = _fits.read_image(p)
        assert np.array_equal(back, arr) and back.shape == arr.shape
        assert hdr["OBJECT"] == "it's a test" and hdr["EXPTIME"] == 53.9 and hdr["FLAG"] is True and hdr["N"] == 7 and np.isnan(hdr["BAD"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_getpsf_gpu_matches_reference_golden(name):
    from pythonphot_b200 import getpsf as gp
    g = load_golden("golden_getpsf.npz")
    out, log = _run(gp, g, name)
    _check(out, g, name)
    assert log.count("too near edge") + log.count("No convergence") == int(g[name + "__nskipped"])


@pytest.mark.gpu
def test_rdpsf_gpu_roundtrip(tmp_path):
    from pythonphot_b200 import getpsf as gp, rdpsf, dao_value
    g = load_golden("golden_getpsf.npz")
    path = os.path.join(str(tmp_path), "psf.fits")
    (gauss, psf, psfmag), _ = _run(gp, g, "r8_f3", psfname=path)
    full, hdr = rdpsf.rdpsf(path)
    assert full.shape == (17, 17) and hdr["NAXIS1"] == 17
    d = np.arange(17) - 8.0
    dx, dy = np.meshgrid(d, d)
    want = dao_value.dao_value(dx.ravel(), dy.ravel(), gauss, psf).reshape(17, 17)
    assert np.array_equal(full, want)
    assert np.unravel_index(np.argmax(full), full.shape) == (8, 8)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["past_end", "above_first_row", "both"])
def test_rinter_gpu_flat_index_semantics_at_the_table_edges(case):
    """rinter.py indexes the flattened table: rows wrap into each other, indices past the end read the padded last
    element, negative ones wrap numpy-style.  Golden: the unmodified reference on points in and around a 24x24 table
    (generated with oracle/refloader.py, see tests/golden/golden_rinter_edge.npz)."""
    from pythonphot_b200 import rinter
    g = load_golden("golden_rinter_edge.npz")
    z, dx, dy = rinter.rinter(g["table"], g[case + "__x"], g[case + "__y"], deriv=True)
    assert_close(z, g[case + "__z"], rtol=1e-9, atol=1e-12, what=case + " z")
    assert_close(dx, g[case + "__dfdx"], rtol=1e-9, atol=1e-12, what=case + " dfdx")
    assert_close(dy, g[case + "__dfdy"], rtol=1e-9, atol=1e-12, what=case + " dfdy")
