"""Pins the CPU oracle (oracle/bphot_oracle.c) to the golden vectors produced by the
UNMODIFIED reference (tests/golden/gen_golden.py).  CPU only."""
import numpy as np
import pytest

from oracle import oracle
from tests.conftest import assert_close, load_golden
from tests.golden.cases import APER_CASES


def test_pixwt_docstring_value():
    # pixwt.py:25-30: Pixwt(3.23,4.22,3.44,5,7) ==> 0.6502
    w = oracle.Pixwt(3.23, 4.22, 3.44, [5.0], [7.0])
    assert abs(w[0] - 0.6502) < 5e-5


def test_pixwt_golden():
    g = load_golden("golden_pixwt.npz")
    for i in range(g["xc"].size):
        m = g["pid"] == i
        w = oracle.Pixwt(g["xc"][i], g["yc"][i], g["r"][i], g["x"][m], g["y"][m])
        assert_close(w, g["w"][m], rtol=1e-12, atol=1e-13, what="pixwt case %d" % i)
    w = oracle.Pixwt(3.5, 3.5, 2.0, g["half_x"], g["half_y"])
    assert_close(w, g["half_w"], rtol=1e-12, atol=1e-13, what="pixwt half-integer centre")
    assert_close(oracle.Pixwt(3.23, 4.22, 3.44, [5.0], [7.0]), g["doc"], rtol=1e-13)


def test_mmm_golden():
    g = load_golden("golden_mmm.npz")
    off = 0
    for n, out in zip(g["lens"], g["out"]):
        v = g["vec"][off:off + n]
        off += n
        got = oracle.mmm(v, readnoise=out[3], minsky=int(out[4]))
        assert_close(got, out[:3], rtol=1e-12, atol=1e-12, what="mmm n=%d" % n)


def test_psfmodel_golden():
    g = load_golden("golden_psfmodel.npz")
    f, pder = oracle.daoerf(g["xx"], g["yy"], g["gauss"])
    assert_close(f, g["f"], rtol=1e-12, atol=1e-300, what="daoerf f")
    assert_close(pder, g["pder"], rtol=1e-11, atol=1e-12, what="daoerf pder")
    half = (g["psf"].shape[0] - 1) / 2.0
    z, dx, dy = oracle.rinter(g["psf"], 2 * g["xx"] + half, 2 * g["yy"] + half)
    assert_close(z, g["z"], rtol=1e-12, atol=1e-12, what="rinter z")
    assert_close(dx, g["dfdx"], rtol=1e-12, atol=1e-12, what="rinter dfdx")
    assert_close(dy, g["dfdy"], rtol=1e-12, atol=1e-12, what="rinter dfdy")
    v, dvx, dvy = oracle.dao_value(g["xx"], g["yy"], g["gauss"], g["psf"])
    assert_close(v, g["v"], rtol=1e-12, atol=1e-12, what="dao_value")
    assert_close(dvx, g["dvx"], rtol=1e-11, atol=1e-11, what="dao_value dvdx")
    assert_close(dvy, g["dvy"], rtol=1e-11, atol=1e-11, what="dao_value dvdy")
    ve, _, _ = oracle.dao_value(g["xe"], g["yy"][:10], g["gauss"], g["psf"])
    assert np.all(ve == 0.0) and np.all(g["ve"] == 0.0)


@pytest.mark.parametrize("mode", ["py3", "py2"])
@pytest.mark.parametrize("cname", [c[0] for c in APER_CASES])
def test_aper_golden(mode, cname):
    g = load_golden("golden_aper.npz")
    kw = dict(dict(APER_CASES)[cname])
    if mode == "py2" and not kw.get("exact"):
        pytest.skip("py2 == py3 for exact=False")
    r = oracle.aper(g["image"], g["x"], g["y"], phpadu=float(g["phpadu"]), zeropoint=25.0, compat=mode, **kw)
    names = ("mag", "magerr", "flux", "fluxerr", "sky", "skyerr", "badflag")
    for nm, a in zip(names, r):
        want = g["%s__%s__%s" % (mode, cname, nm)]
        if nm == "badflag":
            assert np.array_equal(a, want), "badflag %s" % cname
        else:
            # flux is a difference of ~1e4-scale sums: allow 1e-9 of the summed magnitude via atol
            assert_close(a, want, rtol=1e-9, atol=1e-7 if nm in ("flux", "fluxerr") else 1e-10,
                         what="%s %s %s" % (mode, cname, nm))


PK_COLS = {"norecent": ("errmag", "chi", "sharp", "niter", "scale", "chi2"),
           "noise": ("errmag", "chi", "sharp", "niter", "scale", "x", "y"),
           "pkfit_rec": ("errmag", "chi", "sharp", "niter", "scale", "x", "y"),
           "pkfit_norec": ("errmag", "chi", "sharp", "niter", "scale", "x", "y")}


@pytest.mark.parametrize("radius", [3.0, 5.0])
@pytest.mark.parametrize("kind", list(PK_COLS))
def test_pkfit_golden(kind, radius):
    g = load_golden("golden_pkfit.npz")
    want = g["%s__r%d" % (kind, radius)]
    variant = {"norecent": oracle.PKFIT_NORECENT, "noise": oracle.PKFIT_NOISE,
               "pkfit_rec": oracle.PKFIT, "pkfit_norec": oracle.PKFIT}[kind]
    o = oracle.pkfit(variant, g["image"], g["gauss"], g["psf"], float(g["ronois"]), float(g["phpadu"]),
                     g["noise"], g["mask"], g["scale0"], g["x"], g["y"], g["sky"], radius,
                     recenter=(kind != "pkfit_norec"))
    cols = PK_COLS[kind]
    assert np.array_equal(o["niter"], want[:, 3].astype(np.int32)), \
        "niter differs at %s" % np.nonzero(o["niter"] != want[:, 3])[0][:8]
    for j, nm in enumerate(cols):
        if nm == "niter":
            continue
        w = want[:, j]
        got = o[nm]
        if nm in ("x", "y", "chi2"):
            # early-return tuples of the reference drop these columns (padded NaN in the golden file)
            keep = ~np.isnan(w)
            got, w = got[keep], w[keep]
        if nm == "sharp":
            # Off-frame stars see only the far Gaussian tail, where Phi(u2)-Phi(u1) cancels to ~1e-7 of an
            # erf ulp: scipy's erf (reference) and libm's erf (oracle) then differ by ~1e-9 relative in the
            # model itself.  Third-party ulp-level arithmetic, so those stars get 1e-6.
            ny_, nx_ = g["image"].shape
            off = ~((g["x"] >= 0) & (g["x"] <= nx_ - 1) & (g["y"] >= 0) & (g["y"] <= ny_ - 1))
            assert_close(got[off], w[off], rtol=1e-6, atol=1e-12, what="%s r=%g sharp(off-frame)" % (kind, radius))
            got, w = got[~off], w[~off]
        assert_close(got, w, rtol=1e-9, atol=1e-12, what="%s r=%g %s" % (kind, radius, nm))


@pytest.mark.parametrize("mode", ["py3", "py2"])
def test_fragile_apertures_are_checked_against_the_admissible_outcomes(mode):
    """oracle/compare.py on the reference's own goldens: every aperture is compared -- robust ones one to one, fragile ones
    (a bad pixel where Pixwt returns 0 +- 1e-16) against the set of outcomes the reference's arithmetic admits.  The
    'every fragile pixel left out' outcome (what a geometric classification of the pixels yields) must pass too, and a
    wrong value on a robust or on a fragile aperture must not."""
    from oracle import compare
    g = load_golden("golden_aper.npz")
    kw = dict(dict(APER_CASES)["exact_4050"])
    o = oracle.aper(g["image"], g["x"], g["y"], phpadu=float(g["phpadu"]), zeropoint=25.0, compat=mode, return_extra=True, **kw)
    want = {nm: g["%s__exact_4050__%s" % (mode, nm)] for nm in compare.APER_NAMES}
    fragile, alt = o[9], o[10]
    assert fragile.sum() > 20
    rep = compare.compare_aper(dict(zip(compare.APER_NAMES, o[:7])), want, fragile, o[7], o[7], o[8], o[8], alt=alt)
    assert rep["mismatches"] == 0 and rep["fragile_checked"] == rep["fragile"] == int(fragile.sum())
    # an implementation that leaves every fragile pixel out
    other = {nm: np.array(a, copy=True) for nm, a in zip(compare.APER_NAMES, o[:7])}
    for nm in ("flux", "fluxerr", "mag", "magerr", "badflag"):
        other[nm][fragile] = alt[nm][fragile]
    assert (other["badflag"] != o[6]).any(), "the two outcomes should differ somewhere"
    assert compare.compare_aper(other, want, fragile, alt=alt)["mismatches"] == 0
    # wrong values are caught on both kinds of aperture
    for sel in (fragile, ~fragile & np.isfinite(o[2])):
        i, k = np.argwhere(sel)[3]
        broken = {nm: np.array(a, copy=True) for nm, a in other.items()}
        broken["flux"][i, k] *= 1.0 + 1e-6
        assert compare.compare_aper(broken, want, fragile, alt=alt)["mismatches"] == 1
    # a flag raised on a fragile aperture whose fragile pixels are all non-finite needs the zeroed sum with it
    only_nan = fragile & ((alt["kind"] & 4) == 0) & (alt["badflag"] == 0)
    if only_nan.any():
        i, k = np.argwhere(only_nan)[0]
        broken = {nm: np.array(a, copy=True) for nm, a in other.items()}
        broken["badflag"][i, k] = 1
        assert compare.compare_aper(broken, want, fragile, alt=alt)["mismatches"] == 1


@pytest.mark.parametrize("radius", [3.0, 5.0])
def test_pkfit_norecenter_golden(radius):
    """The fourth fit variant, pkfit_norecenter.pkfit_norecenter (pkfit_norecenter.py:346-657), against the reference run
    under its Python-2 semantics (tests/golden/gen_golden_norecenter.py)."""
    from oracle import compare
    g = load_golden("golden_pkfit.npz")
    w = load_golden("golden_pkfit_norecenter.npz")["norecenter__r%d" % radius]
    cols = ("errmag", "chi", "sharp", "niter", "scale")
    want = {c: w[:, j] for j, c in enumerate(cols)}
    o = oracle.pkfit(oracle.PKFIT_NORECENTER, g["image"], g["gauss"], g["psf"], float(g["ronois"]), float(g["phpadu"]), None, None,
                     g["scale0"], g["x"], g["y"], g["sky"], radius)
    rep = compare.compare_pkfit(o, want, cols, g["x"], g["y"], g["image"].shape, radius)
    assert rep["mismatches"] == 0 and rep["rank_deficient"] <= 3, rep
    # it is NOT pkfit(recenter=False): the two differ where pixels are rejected or an early return is taken
    p = oracle.pkfit(oracle.PKFIT, g["image"], g["gauss"], g["psf"], float(g["ronois"]), float(g["phpadu"]), None, None,
                     g["scale0"], g["x"], g["y"], g["sky"], radius, recenter=False)
    assert (p["niter"] != o["niter"]).any()


# ----------------------------------------------------------------------------- large annuli / long sky vectors
@pytest.mark.parametrize("mode,cname", [("py3", "r100_180"), ("py3", "r60_260"), ("py2", "r100_180")])
def test_oracle_big_annulus_matches_reference(mode, cname):
    """tests/golden/gen_golden_bigannulus.py: aper with 50 000 - 180 000 annulus pixels per star, reference vs oracle."""
    import hashlib
    from oracle import compare, oracle
    from tests.golden import gen_golden_bigannulus as G
    g = load_golden("golden_aper_bigannulus.npz")
    f = G.field()
    assert hashlib.sha256(np.ascontiguousarray(f["image"]).tobytes()).hexdigest() == str(g["image_sha256"]), "generator drift"
    o = oracle.aper(f["image"], g["x"], g["y"], phpadu=float(g["phpadu"]), zeropoint=25.0, compat=mode, return_extra=True, **G.CASES[cname])
    want = {nm: g["%s__%s__%s" % (mode, cname, nm)] for nm in G.NAMES}
    rep = compare.compare_aper(dict(zip(G.NAMES, o[:7])), want, o[9], alt=o[10])
    assert rep["mismatches"] == 0 and rep["fragile_checked"] == rep["fragile"], rep
    assert int(np.max(o[7])) > 32768


def test_oracle_mmm_long_vectors_match_reference():
    from oracle import oracle
    from tests.golden import gen_golden_bigannulus as G
    g = load_golden("golden_aper_bigannulus.npz")
    for v, want in zip(G.big_vectors(), g["mmm_big"]):
        assert_close(oracle.mmm(v), want, rtol=1e-12, what="mmm n=%d" % v.size)
