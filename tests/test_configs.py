"""BASELINE.json configs[0] and configs[1] at their stated size, against goldens produced by the UNMODIFIED reference
(tests/golden/gen_golden_configs.py; README Examples A and B, /root/reference/README.md:26-82).

configs[0]: find-detected stars of a 2048 x 1489 frame -> aper(exact=True, apr=5, skyrad=[40,50]) in both readings of
aper.py:246-248.  configs[1]: getpsf(psfrad=8, fitrad=3) + pkfit_norecent_noise(1, x, y, sky, 5) on ~1.7 k stars x 3
frames with noise and mask images.  The frames are rebuilt from their seeds and checked against the stored SHA-256.
CPU tests pin the oracle to these goldens; `-m gpu` tests run the product path."""
import hashlib

import numpy as np
import pytest

from tests.conftest import assert_close, load_golden
from tests.golden import gen_golden_configs as G

APER_KW = dict(apr=5.0, skyrad=[40, 50], badpix=[-12000, 60000], exact=True)
FIT_COLS = ("errmag", "chi", "sharp", "niter", "scale")


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def c1():
    g = load_golden("golden_config1.npz")
    f = G.config1_field()
    assert _sha(f["image"]) == str(g["image_sha256"]), "the synthetic generator no longer reproduces the golden frame"
    return g, f


@pytest.fixture(scope="module")
def c2():
    g = load_golden("golden_config2.npz")
    fields = {}
    for band, seed in zip("gri", G.C2_SEEDS):
        f = G.config2_field(seed)
        for k in ("image", "noise", "mask"):
            assert _sha(f[k]) == str(g["%s__%s_sha256" % (band, k)]), "generator drift: %s %s" % (band, k)
        fields[band] = f
    return g, fields


def _aper_check(got, g, mode, o):
    from oracle import compare
    want = {nm: g["%s__%s" % (mode, nm)] for nm in compare.APER_NAMES}
    rep = compare.compare_aper(got, want, o[9], alt=o[10])
    assert rep["mismatches"] == 0 and rep["fragile_checked"] == rep["fragile"], rep
    return rep


# ----------------------------------------------------------------------------- CPU: the oracle at config size
@pytest.mark.parametrize("mode", ["py3", "py2"])
def test_config1_oracle_matches_reference(c1, mode):
    from oracle import compare, oracle
    g, f = c1
    x, y = g["find__x"], g["find__y"]
    assert x.size > 2000
    o = oracle.aper(f["image"], x, y, phpadu=float(g["phpadu"]), zeropoint=25.0, compat=mode, nthreads=0, return_extra=True, **APER_KW)
    rep = _aper_check(dict(zip(compare.APER_NAMES, o[:7])), g, mode, o)
    assert rep["stars"] == x.size


def test_config1_find_oracle_matches_reference(c1):
    from oracle import find_oracle
    g, f = c1
    want = [g["find__" + nm] for nm in ("x", "y", "flux", "sharp", "round")]
    got = find_oracle.find(f["image"].astype(np.float64), G.C1["hmin"], G.C1["fwhm"], G.C1["roundlim"], G.C1["sharplim"])
    assert got[0].size == want[0].size
    for a, b, nm in zip(got, want, ("x", "y", "flux", "sharp", "round")):
        assert_close(a, b, rtol=1e-9, atol=1e-9, what="find " + nm)


@pytest.mark.parametrize("band", list("gri"))
def test_config2_oracle_fits_match_reference(c2, band):
    from oracle import compare, oracle
    g, fields = c2
    f = fields[band]
    x, y, sky = g[band + "__x"], g[band + "__y"], g[band + "__sky"]
    want = {c: g[band + "__fit"][:, j] for j, c in enumerate(FIT_COLS)}
    o = oracle.pkfit(oracle.PKFIT_NORECENT, f["image"], g[band + "__gauss"], g[band + "__psf"], float(g[band + "__ronois"]),
                     float(g[band + "__phpadu"]), f["noise"], f["mask"], np.ones(x.size), x, y, sky, 5.0, nthreads=0)
    rep = compare.compare_pkfit(o, want, FIT_COLS, x, y, f["image"].shape, 5.0)
    assert rep["mismatches"] == 0 and rep["rank_deficient"] < 0.01 * x.size, rep


# ----------------------------------------------------------------------------- GPU: the product path at config size
@pytest.mark.gpu
def test_config1_gpu_find_then_aper(c1):
    """README Example A end to end on the GPU: the detected star list equals the reference's, and aper at those stars
    equals the reference's in both compat modes (every aperture checked, fragile ones against the admissible outcomes)."""
    import pythonphot_b200 as pp
    from oracle import compare, oracle
    g, f = c1
    r = pp.find.find(f["image"], G.C1["hmin"], G.C1["fwhm"], G.C1["roundlim"], G.C1["sharplim"], verbose=False)
    for a, nm in zip(r, ("x", "y", "flux", "sharp", "round")):
        assert_close(a, g["find__" + nm], rtol=1e-9, atol=1e-9, what="find " + nm)
    x, y = np.asarray(r[0]), np.asarray(r[1])
    for mode in ("py3", "py2"):
        a = pp.aper.aper(f["image"], x, y, phpadu=float(g["phpadu"]), zeropoint=25.0, verbose=False, compat=mode, **APER_KW)
        got = {nm: np.asarray(v, dtype=np.float64) for nm, v in zip(compare.APER_NAMES, a[:7])}
        o = oracle.aper(f["image"], g["find__x"], g["find__y"], phpadu=float(g["phpadu"]), zeropoint=25.0, compat=mode, nthreads=0,
                        return_extra=True, **APER_KW)
        _aper_check(got, g, mode, o)


@pytest.mark.gpu
@pytest.mark.parametrize("band", list("gri"))
def test_config2_gpu_getpsf_then_forced_fits(c2, band):
    """README Example B on the GPU: find -> aper -> getpsf(psfrad=8, fitrad=3) reproduce the reference's PSF model, and
    pkfit_norecent_noise(1, x, y, sky, 5) over every detected star reproduces its fits (niter exact, values 1e-9)."""
    import pythonphot_b200 as pp
    from oracle import compare
    g, fields = c2
    f = fields[band]
    r = pp.find.find(f["image"], G.C2["hmin"], G.C2["fwhm"], [-1.0, 1.0], [0.2, 1.0], verbose=False)
    assert_close(r[0], g[band + "__x"], rtol=1e-9, atol=1e-9, what="find x")
    assert_close(r[1], g[band + "__y"], rtol=1e-9, atol=1e-9, what="find y")
    x, y = g[band + "__x"], g[band + "__y"]
    ph, ron = float(g[band + "__phpadu"]), float(g[band + "__ronois"])
    a = pp.aper.aper(f["image"], x, y, phpadu=ph, apr=5.0, zeropoint=25.0, skyrad=[40, 50], badpix=[-12000, 60000], exact=False,
                     verbose=False)
    mag, sky = np.asarray(a[0], float).reshape(-1), np.asarray(a[4], float).reshape(-1)
    mag = np.where(np.isfinite(mag), mag, 99.0)
    assert_close(mag, g[band + "__mag"], rtol=1e-9, atol=1e-9, what="aper mag")
    assert_close(sky, g[band + "__sky"], rtol=1e-9, atol=1e-9, what="aper sky")
    gauss, psf, psfmag = pp.getpsf.getpsf(f["image"], x, y, g[band + "__mag"], g[band + "__sky"], ron, ph, g[band + "__idpsf"], 8, 3,
                                          None, verbose=False)
    assert_close(gauss, g[band + "__gauss"], rtol=1e-9, what="gauss")
    want_psf = g[band + "__psf"]
    assert_close(psf, want_psf, rtol=1e-9, atol=1e-9 * np.abs(want_psf).max(), what="psf table")
    assert_close(psfmag, g[band + "__psfmag"], rtol=1e-9, what="psfmag")
    pk = pp.pkfit_norecent_noise.pkfit_class(f["image"], g[band + "__gauss"], g[band + "__psf"], ron, ph, f["noise"], f["mask"])
    res = pk.pkfit_norecent_noise(np.ones(x.size), x, y, g[band + "__sky"], 5)
    got = dict(zip(FIT_COLS, res))
    want = {c: g[band + "__fit"][:, j] for j, c in enumerate(FIT_COLS)}
    rep = compare.compare_pkfit(got, want, FIT_COLS, x, y, f["image"].shape, 5.0)
    assert rep["mismatches"] == 0 and rep["rank_deficient"] < 0.01 * x.size, rep
