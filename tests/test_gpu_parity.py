"""Parity of the CUDA path (through libbphot's C ABI) against the golden vectors of the unmodified
reference and against the CPU oracle on seeded synthetic fields.  Needs a B200: `pytest -m gpu`.

Tolerances (north_star): counts / flags / iteration numbers bit-exact; fp64 quantities 1e-9 relative.
Aperture fluxes are differences of ~1e4..1e6-ADU sums, so they additionally get an absolute 1e-7 ADU
(1e-11 of the summed counts)."""
import ctypes as C

import numpy as np
import pytest

from tests.conftest import assert_close, load_golden
from tests.golden.cases import APER_CASES

pytestmark = pytest.mark.gpu

NAMES = ("mag", "magerr", "flux", "fluxerr", "sky", "skyerr", "badflag")


@pytest.fixture(scope="module")
def pp():
    import pythonphot_b200
    pythonphot_b200._lib.torch_cuda()
    return pythonphot_b200


@pytest.fixture(scope="module")
def oracle():
    from oracle import oracle as o
    return o


# ------------------------------------------------------------------ device math vs goldens
def test_pixwt_golden(pp):
    g = load_golden("golden_pixwt.npz")
    for i in range(g["xc"].size):
        m = g["pid"] == i
        w = pp.pixwt.Pixwt(g["xc"][i], g["yc"][i], g["r"][i], g["x"][m], g["y"][m])
        assert_close(w, g["w"][m], rtol=1e-12, atol=1e-12, what="pixwt %d" % i)
    assert_close(pp.pixwt.Pixwt(3.5, 3.5, 2.0, g["half_x"], g["half_y"]), g["half_w"], rtol=1e-12, atol=1e-12)
    assert abs(pp.pixwt.Pixwt(3.23, 4.22, 3.44, 5, 7) - 0.6502) < 5e-5  # pixwt.py:25-30


def test_psfmodel_golden(pp):
    g = load_golden("golden_psfmodel.npz")
    f, pder = pp.daoerf.daoerf(g["xx"], g["yy"], g["gauss"])
    assert_close(f, g["f"], rtol=1e-11, atol=1e-300, what="daoerf f")
    assert_close(pder, g["pder"], rtol=1e-10, atol=1e-11, what="daoerf pder")
    half = (g["psf"].shape[0] - 1) / 2.0
    z, dx, dy = pp.rinter.rinter(g["psf"], 2 * g["xx"] + half, 2 * g["yy"] + half, deriv=True)
    assert_close(z, g["z"], rtol=1e-12, atol=1e-12, what="rinter z")
    assert_close(dx, g["dfdx"], rtol=1e-12, atol=1e-12, what="rinter dfdx")
    assert_close(dy, g["dfdy"], rtol=1e-12, atol=1e-12, what="rinter dfdy")
    v, dvx, dvy = pp.dao_value.dao_value(g["xx"], g["yy"], g["gauss"], g["psf"], deriv=True)
    assert_close(v, g["v"], rtol=1e-11, atol=1e-11, what="dao_value")
    assert_close(dvx, g["dvx"], rtol=1e-10, atol=1e-10, what="dvdx")
    assert_close(dvy, g["dvy"], rtol=1e-10, atol=1e-10, what="dvdy")
    ve = pp.dao_value.dao_value(g["xe"], g["yy"][:10], g["gauss"], g["psf"], deriv=False)
    assert np.all(ve == 0.0)  # dao_value.py:70-84


def test_mmm_golden(pp):
    g = load_golden("golden_mmm.npz")
    off = 0
    for n, out in zip(g["lens"], g["out"]):
        v = g["vec"][off:off + n]
        off += n
        got = pp.mmm.mmm(v, readnoise=out[3], minsky=int(out[4]))
        assert_close(got, out[:3], rtol=1e-10, atol=1e-10, what="mmm n=%d" % n)


def test_mmm_batch_vs_oracle_ties(pp, oracle):
    rng = np.random.default_rng(77)
    vecs = []
    for t in range(400):
        n = int(rng.integers(20, 3000))
        v = np.round(rng.normal(300, rng.uniform(0.5, 8), n)) if t % 2 else rng.normal(300, 9, n)
        v[rng.integers(0, n, n // 10 + 1)] += np.round(rng.exponential(300, n // 10 + 1))
        vecs.append(v.astype(np.float32) if t % 3 == 0 else v)
    vecs32 = [v.astype(np.float32) for v in vecs]
    for batch in (vecs, vecs32):
        got, iters = pp.mmm.mmm_batch(batch, return_iters=True)
        for i, v in enumerate(batch):
            want = oracle.mmm(v.astype(np.float64), return_info=True)
            assert iters[i] == want[3], "mmm iterations differ for vector %d" % i
            assert_close(got[i], want[:3], rtol=1e-10, atol=1e-10, what="mmm vec %d" % i)


# ------------------------------------------------------------------ aper vs the reference's goldens
def _check_aper(got, want, o, what, nsky=None, want_nsky=None, mmm_iter=None):
    """got/want: dicts of arrays; `o` = the oracle's return_extra tuple for the same call.  badflag bit-exact and values
    1e-9 on robust apertures; a fragile aperture (its flag hinges on the sign of a rounding error in the reference itself,
    see oracle/bphot_oracle.c) must equal one of the outcomes the reference admits (oracle/compare.py) -- nothing is skipped."""
    from oracle import compare
    rep = compare.compare_aper(got, want, o[9], nsky, want_nsky, mmm_iter, o[8] if mmm_iter is not None else None, alt=o[10])
    assert rep["mismatches"] == 0, "%s: %r" % (what, rep)
    assert rep["fragile_checked"] == rep["fragile"]
    return rep


@pytest.mark.parametrize("mode", ["py3", "py2"])
@pytest.mark.parametrize("cname", [c[0] for c in APER_CASES])
def test_aper_golden(pp, oracle, mode, cname):
    g = load_golden("golden_aper.npz")
    kw = dict(dict(APER_CASES)[cname])
    if mode == "py2" and not kw.get("exact"):
        pytest.skip("py2 == py3 for exact=False")
    r = pp.aper.aper(g["image"], g["x"], g["y"], phpadu=float(g["phpadu"]), zeropoint=25.0, verbose=False,
                     compat=mode, return_info=True, **kw)
    got = dict(zip(NAMES, r[:7]))
    n = g["x"].size
    for k in ("mag", "magerr", "flux", "fluxerr", "badflag"):
        got[k] = np.asarray(got[k]).reshape(n, -1)
    want = {nm: g["%s__%s__%s" % (mode, cname, nm)] for nm in NAMES}
    o = oracle.aper(g["image"], g["x"], g["y"], phpadu=float(g["phpadu"]), zeropoint=25.0, compat=mode,
                    return_extra=True, **kw)
    rep = _check_aper(got, want, o, "%s/%s" % (mode, cname), r[8]["nsky"], o[7], r[8]["mmm_iter"])
    print("fragile apertures: %d (checked against the admissible outcomes), flag differs from the oracle's on %d" % (
        rep["fragile"], rep["fragile_flag_differs"]))


def test_aper_scalar_star_squeezes(pp):
    g = load_golden("golden_aper.npz")
    r = pp.aper.aper(g["image"], float(g["x"][3]), float(g["y"][3]), phpadu=float(g["phpadu"]), apr=5.0,
                     skyrad=[40, 50], badpix=[-12000, 60000], exact=True, verbose=False)
    assert np.shape(r[0]) == (1,) and np.shape(r[4]) == () and isinstance(r[7], str)  # aper.py:419-427


# ------------------------------------------------------------------ aper vs the oracle, larger fields
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("exact,compat", [(True, "py3"), (True, "py2"), (False, "py3")])
def test_aper_vs_oracle_field(pp, oracle, dtype, exact, compat):
    from pythonphot_b200 import synth
    f = synth.make_field(700, 900, 3000, seed=31, dtype=dtype, with_noise=False, edge_frac=0.03)
    kw = dict(phpadu=f["phpadu"], apr=np.array([3., 5., 8.]), zeropoint=25.0, skyrad=[40, 50],
              badpix=[-12000, 60000], exact=exact)
    r = pp.aper.aper(f["image"], f["x"], f["y"], verbose=False, compat=compat, return_info=True, **kw)
    got = dict(zip(NAMES, r[:7]))
    o = oracle.aper(f["image"], f["x"], f["y"], compat=compat, nthreads=0, return_extra=True, **kw)
    want = dict(zip(NAMES, o[:7]))
    rep = _check_aper(got, want, o, "field %s exact=%s %s" % (np.dtype(dtype).name, exact, compat), r[8]["nsky"], o[7], r[8]["mmm_iter"])
    print("fragile: %d of %d, flag differs from the oracle's on %d" % (rep["fragile"], o[9].size, rep["fragile_flag_differs"]))


@pytest.mark.parametrize("skyrad,fit96", [([12, 19.8], 768), ([40, 51.5], 3072)])
def test_aper_narrow_ctas_and_wide_retry(pp, skyrad, fit96):
    """Stage 1 with 96-thread CTAs forced: stars whose annulus exceeds 96 x K keys are redone by the 128-thread pass
    (default radii never need it).  Every column must equal the default launch bit for bit."""
    import torch
    from pythonphot_b200 import synth
    from pythonphot_b200.aper import aper_device
    f = synth.make_field(600, 700, 2500, seed=77, dtype=np.float32, with_noise=False, edge_frac=0.05)
    img, x, y = (torch.from_numpy(np.ascontiguousarray(f[k])).cuda() for k in ("image", "x", "y"))
    kw = dict(phpadu=f["phpadu"], apr=np.array([3., 5.]), skyrad=skyrad, badpix=[-12000, 60000], exact=True)
    a = aper_device(img, x, y, **kw)
    b = aper_device(img, x, y, _narrow_ctas=True, **kw)
    nsky = a["nsky"].cpu().numpy()
    assert (nsky > fit96).any(), "no star exercises the retry pass"
    if fit96 == 768:
        assert ((nsky > 0) & (nsky <= fit96 - 40)).any(), "no star stays on the narrow pass"
    for k in a:
        ta, tb = a[k].cpu().numpy(), b[k].cpu().numpy()
        assert np.array_equal(ta, tb, equal_nan=ta.dtype.kind == "f"), k


def test_aper_edge_matrix(pp, oracle):
    """Branches of aper.py: off-frame, edge < apr, nsky < minsky, NaN in annulus / aperture, all-NaN."""
    rng = np.random.default_rng(3)
    img = rng.normal(100, 5, (160, 180))
    img[60:70, 60:70] = np.nan           # NaNs inside an aperture
    img[100:104, 20:90] = 70000.0        # saturated streak through annuli
    img[:, 150:] = np.nan                # a whole region of NaN: nsky after NaN drop < minsky
    x = np.array([-5.0, 0.4, 0.6, 2.0, 90.0, 65.0, 64.2, 30.0, 178.6, 179.0, 165.0, 90.0, 12.5, np.nan])
    y = np.array([50.0, 50.0, 50.0, 3.0, 80.0, 65.0, 60.1, 102.0, 80.0, 80.0, 80.0, 158.9, 12.5, 20.0])
    for kw in (dict(apr=np.array([2., 4., 7.]), skyrad=[10, 15], badpix=[-100, 60000], exact=True),
               dict(apr=np.array([2., 4., 7.]), skyrad=[10, 15], badpix=[0, 0], exact=False),
               dict(apr=np.array([3.]), skyrad=[3, 5], badpix=[-100, 60000], exact=True, minsky=60),
               dict(apr=np.array([2., 3.]), skyrad=[10, 15], badpix=[-100, 60000], exact=True, setskyval=99.0)):
        r = pp.aper.aper(img, x, y, phpadu=2.0, verbose=False, return_info=True, **kw)
        o = oracle.aper(img, x, y, phpadu=2.0, return_extra=True, **kw)
        _check_aper(dict(zip(NAMES, r[:7])), dict(zip(NAMES, o[:7])), o, "edge %r" % (kw,), r[8]["nsky"], o[7])


def test_aper_properties_full_size(pp):
    """BASELINE config 3 shape (4096^2, 100k stars): power-of-two scaling is exact in fp64, a star-list
    permutation permutes results bit-for-bit, and a constant image integrates to pi r^2 (Pixwt closure)."""
    from pythonphot_b200 import synth
    f = synth.make_field(4096, 4096, 100000, seed=1003, dtype=np.float32, with_noise=False)
    kw = dict(phpadu=f["phpadu"], apr=np.array([3., 5., 8.]), skyrad=[40, 50], badpix=[-12000, 60000], exact=True,
              verbose=False)
    a = pp.aper.aper(f["image"], f["x"], f["y"], **kw)
    kw2 = dict(kw, badpix=[-24000, 120000])
    b = pp.aper.aper(f["image"] * np.float32(2), f["x"], f["y"], **kw2)
    assert np.array_equal(np.nan_to_num(b[2], nan=-1.0), np.nan_to_num(2 * a[2], nan=-1.0)), "flux not linear"
    assert np.array_equal(np.nan_to_num(b[4], nan=-1.0), np.nan_to_num(2 * a[4], nan=-1.0)), "sky not linear"
    assert np.array_equal(a[6], b[6])
    perm = np.random.default_rng(0).permutation(f["x"].size)
    c = pp.aper.aper(f["image"], f["x"][perm], f["y"][perm], **kw)
    for i in (0, 1, 2, 3, 4, 5, 6):
        assert np.array_equal(np.nan_to_num(c[i], nan=-1.0), np.nan_to_num(a[i][perm], nan=-1.0)), "permutation changed output %d" % i
    const = np.full((512, 512), 7.0, dtype=np.float32)
    xs = np.random.default_rng(1).uniform(60, 450, 500)
    ys = np.random.default_rng(2).uniform(60, 450, 500)
    for compat in ("py3", "py2"):
        d = pp.aper.aper(const, xs, ys, phpadu=1, apr=np.array([2., 3., 5., 8.]), skyrad=[40, 50], exact=True,
                         setskyval=0.0, verbose=False, compat=compat)
        area = np.pi * np.array([2., 3., 5., 8.]) ** 2
        assert np.abs(d[2] / (7.0 * area) - 1).max() < 1e-9


# ------------------------------------------------------------------ pkfit vs the reference's goldens
PK_COLS = {"norecent": ("errmag", "chi", "sharp", "niter", "scale", "chi2"),
           "noise": ("errmag", "chi", "sharp", "niter", "scale", "x", "y"),
           "pkfit_rec": ("errmag", "chi", "sharp", "niter", "scale", "x", "y"),
           "pkfit_norec": ("errmag", "chi", "sharp", "niter", "scale", "x", "y")}


def _run_pk(pp, kind, g_or_f, scale, x, y, sky, radius, image, noise, mask):
    gauss, psf, ron, ph = g_or_f["gauss"], g_or_f["psf"], float(g_or_f["ronois"]), float(g_or_f["phpadu"])
    if kind == "norecent":
        pk = pp.pkfit_norecent_noise.pkfit_class(image, gauss, psf, ron, ph, noise, mask)
        r = pk.pkfit_norecent_noise(scale, x, y, sky, radius, returnchi2=True)
        return dict(zip(PK_COLS[kind], r))
    if kind == "noise":
        pk = pp.pkfit_noise.pkfit_class(image, gauss, psf, ron, ph, noise, mask)
        return dict(zip(PK_COLS[kind], pk.pkfit_noise(scale, x, y, sky, radius, xyout=True)))
    pk = pp.pkfit.pkfit_class(image, gauss, psf, ron, ph)
    return dict(zip(PK_COLS[kind], pk.pkfit(scale, x, y, sky, radius, xyout=True, recenter=(kind == "pkfit_rec"))))


def _underdetermined(x, y, radius, shape):
    """Stars with fewer than 4 in-frame pixels inside the fitting radius: their 3x3 normal matrix is rank
    deficient, and whether LAPACK meets an exactly zero pivot (-> NaN return, pkfit.py:383-406) or a
    1e-17 one (-> garbage step) hinges on the last bit of erf(); neither implementation is 'right'."""
    ny_, nx_ = shape
    R = int(np.ceil(radius)) + 1
    xi, yi = np.floor(np.nan_to_num(x)).astype(np.int64), np.floor(np.nan_to_num(y)).astype(np.int64)
    cnt = np.zeros(x.size, dtype=np.int64)
    for oy in range(-R, R + 1):
        for ox in range(-R, R + 1):
            px, py = xi + ox, yi + oy
            inside = (px >= 0) & (px < nx_) & (py >= 0) & (py < ny_)
            cnt += inside & (((px - x) ** 2 + (py - y) ** 2) < radius ** 2)
    return (cnt < 4) & np.isfinite(x) & np.isfinite(y)


def _check_pk(got, want, cols, x, y, shape, what, skip=None):
    """want: dict of arrays (NaN = column dropped by a reference early return, only for x/y/chi2)."""
    if skip is not None and skip.any():
        keep_ = ~skip
        got = {k: np.asarray(v)[keep_] for k, v in got.items()}
        want = {k: np.asarray(v)[keep_] for k, v in want.items()}
        x, y = x[keep_], y[keep_]
    assert np.array_equal(got["niter"], want["niter"].astype(np.int32)), \
        "%s: niter differs at %s" % (what, np.nonzero(got["niter"] != want["niter"])[0][:8])
    ny_, nx_ = shape
    off = ~((x >= 0) & (x <= nx_ - 1) & (y >= 0) & (y <= ny_ - 1))
    for nm in cols:
        if nm == "niter":
            continue
        a, w = got[nm], want[nm]
        keep = np.ones(a.size, bool)
        if nm in ("x", "y", "chi2"):
            keep = ~np.isnan(w)
        if nm == "sharp":
            # off-frame stars see only the far Gaussian tail: erf implementations differ at ~1e-9 there
            assert_close(a[keep & off], w[keep & off], rtol=1e-6, atol=1e-12, what=what + " sharp(off-frame)")
            keep = keep & ~off
        assert_close(a[keep], w[keep], rtol=1e-9, atol=1e-12, what="%s %s" % (what, nm))


@pytest.mark.parametrize("radius", [3.0, 5.0])
@pytest.mark.parametrize("kind", list(PK_COLS))
def test_pkfit_golden(pp, kind, radius):
    g = load_golden("golden_pkfit.npz")
    want_rows = g["%s__r%d" % (kind, radius)]
    want = {c: want_rows[:, j] for j, c in enumerate(PK_COLS[kind])}
    got = _run_pk(pp, kind, g, g["scale0"], g["x"], g["y"], g["sky"], radius, g["image"], g["noise"], g["mask"])
    _check_pk(got, want, PK_COLS[kind], g["x"], g["y"], g["image"].shape, "%s r=%g" % (kind, radius))


def test_pkfit_scalar_call_returns_scalars(pp):
    g = load_golden("golden_pkfit.npz")
    pk = pp.pkfit_class(g["image"], g["gauss"], g["psf"], float(g["ronois"]), float(g["phpadu"]), g["noise"], g["mask"])
    r = pk.pkfit_norecent_noise(float(g["scale0"][0]), float(g["x"][0]), float(g["y"][0]), float(g["sky"][0]), 3.0)
    assert len(r) == 5 and isinstance(r[3], int) and isinstance(r[4], float)
    want = g["norecent__r3"][0]
    assert r[3] == int(want[3]) and abs(r[4] - want[4]) <= 1e-9 * abs(want[4])
    r2 = pk.pkfit_noise(float(g["scale0"][0]), float(g["x"][0]), float(g["y"][0]), float(g["sky"][0]), 3.0, xyout=True)
    assert len(r2) == 7
    r3 = pk.pkfit(float(g["scale0"][0]), float(g["x"][0]), float(g["y"][0]), float(g["sky"][0]), 3.0)
    assert len(r3) == 5


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("kind,radius", [("norecent", 5.0), ("noise", 5.0), ("pkfit_rec", 3.0), ("norecent", 2.2)])
def test_pkfit_vs_oracle_field(pp, oracle, dtype, kind, radius):
    from pythonphot_b200 import synth
    f = synth.make_field(600, 800, 4000, seed=41, dtype=dtype, edge_frac=0.03, margin=10.0)
    n = f["x"].size
    scale0 = f["scale"] * np.where(np.arange(n) % 2 == 0, 1.0, 0.5)
    sky = np.full(n, f["sky"])
    x, y = f["xfit"].copy(), f["yfit"].copy()
    x[7] = np.nan
    got = _run_pk(pp, kind, f, scale0, x, y, sky, radius, f["image"], f["noise"], f["mask"])
    variant = {"norecent": oracle.PKFIT_NORECENT, "noise": oracle.PKFIT_NOISE, "pkfit_rec": oracle.PKFIT}[kind]
    want = oracle.pkfit(variant, f["image"], f["gauss"], f["psf"], f["ronois"], f["phpadu"], f["noise"], f["mask"],
                        scale0, x, y, sky, radius, recenter=True, nthreads=0)
    skip = _underdetermined(x, y, radius, f["image"].shape)
    assert skip.sum() < 0.01 * n
    _check_pk(got, want, PK_COLS[kind], x, y, f["image"].shape, "field %s %s r=%g" % (np.dtype(dtype).name, kind, radius), skip)


def test_pkfit_properties_full_size(pp):
    """BASELINE config 4 shape (8192^2, 1M stars, forced fits, radius 5): a permutation of the star list
    permutes the results bit-for-bit; converged fits recover the planted scale; niter within [1, 25]."""
    from pythonphot_b200 import synth
    f = synth.make_field(8192, 8192, 1000000, seed=1004, dtype=np.float32)
    n = f["x"].size
    pk = pp.pkfit_class(f["image"], f["gauss"], f["psf"], f["ronois"], f["phpadu"], f["noise"], f["mask"])
    sky = np.full(n, f["sky"])
    one = np.ones(n)
    a = pk.pkfit_norecent_noise(one, f["xfit"], f["yfit"], sky, 5.0)
    perm = np.random.default_rng(0).permutation(n)
    b = pk.pkfit_norecent_noise(one, f["xfit"][perm], f["yfit"][perm], sky, 5.0)
    for i in range(5):
        assert np.array_equal(np.nan_to_num(b[i], nan=-1.0), np.nan_to_num(a[i][perm], nan=-1.0))
    niter = a[3]
    assert niter.min() >= 0 and niter.max() <= 25
    # one star per 8x8 px: the field is heavily blended, so the planted scale is recovered only roughly
    # (the oracle gives a median error of 0.19 for scale > 50 on a 20k-star sample of this exposure)
    fin = np.isfinite(a[4])
    good = fin & (f["scale"] > 50)
    rel = np.abs(a[4][good] / f["scale"][good] - 1)
    assert fin.mean() > 0.85 and good.sum() > 0.2 * n and np.median(rel) < 0.3


# ------------------------------------------------------------------ the host-buffer C ABI entry points
def test_c_abi_host_entry_points(pp, oracle):
    from pythonphot_b200 import _lib, synth
    f = synth.make_field(300, 320, 200, seed=5, dtype=np.float32, margin=55.0)
    lib = _lib.lib()
    n, apr = f["x"].size, np.array([3.0, 6.0])
    na = apr.size
    outs = [np.empty((n, na)) for _ in range(4)]
    sky, skyerr = np.empty(n), np.empty(n)
    bad = np.empty((n, na), np.uint8)
    nsky, it = np.empty(n, np.int32), np.empty(n, np.int32)
    img = np.ascontiguousarray(f["image"])
    p = _lib.nptr
    rc = lib.bphot_aper_host(p(img), _lib.F32, img.shape[0], img.shape[1], p(f["x"]), p(f["y"]), n, f["phpadu"], p(apr), na,
                             25.0, 40.0, 50.0, -12000.0, 60000.0, 0, None, 20, 1, 0.0, 0,
                             p(outs[0]), p(outs[1]), p(outs[2]), p(outs[3]), p(sky), p(skyerr), p(bad), p(nsky), p(it), 0)
    _lib.check(rc)
    o = oracle.aper(img, f["x"], f["y"], phpadu=f["phpadu"], apr=apr, skyrad=(40, 50), badpix=(-12000, 60000), exact=True,
                    return_extra=True)
    got = dict(mag=outs[0], magerr=outs[1], flux=outs[2], fluxerr=outs[3], sky=sky, skyerr=skyerr, badflag=bad)
    _check_aper(got, dict(zip(NAMES, o[:7])), o, "bphot_aper_host", nsky, o[7])

    res = {k: np.empty(n) for k in ("errmag", "chi", "sharp", "scale", "x", "y", "chi2")}
    niter = np.empty(n, np.int32)
    nz, mk = np.ascontiguousarray(f["noise"]), np.ascontiguousarray(f["mask"])
    sc, skyv = np.ones(n), np.full(n, f["sky"])
    rc = lib.bphot_pkfit_host(_lib.PKFIT_NORECENT, p(img), _lib.F32, p(nz), _lib.F32, p(mk), img.shape[0], img.shape[1],
                              p(f["gauss"]), p(f["psf"]), f["psf"].shape[1], f["ronois"], f["phpadu"],
                              p(sc), p(f["xfit"]), p(f["yfit"]), p(skyv), 4.0, 25, 0, n,
                              p(res["errmag"]), p(res["chi"]), p(res["sharp"]), p(niter), p(res["scale"]), p(res["x"]),
                              p(res["y"]), p(res["chi2"]), 0)
    _lib.check(rc)
    res["niter"] = niter
    want = oracle.pkfit(oracle.PKFIT_NORECENT, img, f["gauss"], f["psf"], f["ronois"], f["phpadu"], nz, mk, sc, f["xfit"],
                        f["yfit"], skyv, 4.0)
    _check_pk(res, want, PK_COLS["norecent"], f["xfit"], f["yfit"], img.shape, "bphot_pkfit_host")


def test_c_abi_errors_are_reported(pp):
    from pythonphot_b200 import _lib
    lib = _lib.lib()
    rc = lib.bphot_aper_device(None, 0, 10, 10, 10, None, None, 1, 1.0, None, 1, 25.0, 40.0, 50.0, 0.0, 0.0, 0, None,
                               20, 1, 0.0, 0, None, None, None, None, None, None, None, None, None, None)
    assert rc == -1 and b"null" in lib.bphot_last_error()
    # an annulus too large for one CTA's shared memory is not an error (global-memory sort path): here it lies off the frame
    r = pp.aper.aper(np.zeros((64, 64), np.float32), [30.0], [30.0], skyrad=[100, 400], verbose=False)
    assert float(np.asarray(r[4]).reshape(-1)[0]) == 0.0 and float(np.asarray(r[5]).reshape(-1)[0]) == 0.0  # aper.py:269-272


# ------------------------------------------------------------------ annuli above one CTA's shared memory
@pytest.mark.parametrize("mode,cname", [("py3", "r100_180"), ("py3", "r60_260"), ("py2", "r100_180")])
def test_aper_big_annulus_golden(pp, oracle, mode, cname):
    """skyrad=[100,180] / [60,260] (70 000 - 180 000 annulus pixels per star; one CTA sorts at most 32 768): rows in global
    memory + segmented device sort + warp-per-star mmm, against the UNMODIFIED reference (tests/golden/gen_golden_bigannulus.py),
    nsky / mmm iterations against the oracle.  fp32 image, then the same pixels as fp64."""
    import hashlib
    from tests.golden import gen_golden_bigannulus as G
    g = load_golden("golden_aper_bigannulus.npz")
    f = G.field()
    assert hashlib.sha256(np.ascontiguousarray(f["image"]).tobytes()).hexdigest() == str(g["image_sha256"]), "generator drift"
    kw = G.CASES[cname]
    n = g["x"].size
    want = {nm: g["%s__%s__%s" % (mode, cname, nm)] for nm in NAMES}
    o = oracle.aper(f["image"], g["x"], g["y"], phpadu=float(g["phpadu"]), zeropoint=25.0, compat=mode, return_extra=True, **kw)
    for img in (f["image"], f["image"].astype(np.float64)):
        r = pp.aper.aper(img, g["x"], g["y"], phpadu=float(g["phpadu"]), zeropoint=25.0, verbose=False, compat=mode,
                         return_info=True, **kw)
        got = dict(zip(NAMES, r[:7]))
        for k in ("mag", "magerr", "flux", "fluxerr", "badflag"):
            got[k] = np.asarray(got[k]).reshape(n, -1)
        _check_aper(got, want, o, "%s/%s/%s" % (mode, cname, img.dtype), r[8]["nsky"], o[7], r[8]["mmm_iter"])
    assert int(np.max(o[7])) > 32768


def test_mmm_big_vectors_golden(pp, oracle):
    """mmm.mmm on vectors longer than one CTA sorts (40 000 - 262 144 elements, NaNs inside) against the reference's own
    outputs; as one batch together with short vectors (the longest decides the path), fp64 and fp32, iterations against the oracle."""
    from tests.golden import gen_golden_bigannulus as G
    g = load_golden("golden_aper_bigannulus.npz")
    vecs = G.big_vectors()
    for v, want in zip(vecs, g["mmm_big"]):
        assert_close(pp.mmm.mmm(v), want, rtol=1e-9, atol=1e-12, what="mmm n=%d" % v.size)
    rng = np.random.default_rng(9)
    mixed = [vecs[0], rng.normal(300, 5, 500), vecs[2], rng.normal(10, 1, 19), np.full(100, np.nan)]
    for batch in (mixed, [v.astype(np.float32) for v in mixed]):
        got, iters = pp.mmm.mmm_batch(batch, return_iters=True)
        for i, v in enumerate(batch):
            want = oracle.mmm(v.astype(np.float64), return_info=True)
            assert iters[i] == want[3], "mmm iterations differ for vector %d" % i
            assert_close(got[i], want[:3], rtol=1e-9, atol=1e-12, what="mmm vec %d" % i)


def test_aper_big_annulus_many_stars_and_sigclip(pp, oracle):
    """The same path with more stars than one batch of rows holds (2^27 keys per batch / ~96 000 slots per star = 1 390
    stars), and with skyalgorithm='sigmaclipping' (aper.py:275-282) on a subset against oracle/sigclip_oracle.py."""
    from oracle import sigclip_oracle
    from pythonphot_b200 import synth
    f = synth.make_field(700, 700, 1500, seed=515, dtype=np.float32, with_noise=False)
    kw = dict(apr=np.array([5.0]), skyrad=[100, 200], badpix=[-12000, 60000], exact=False)
    n = f["x"].size
    r = pp.aper.aper(f["image"], f["x"], f["y"], phpadu=f["phpadu"], zeropoint=25.0, verbose=False, return_info=True, **kw)
    o = oracle.aper(f["image"], f["x"], f["y"], phpadu=f["phpadu"], zeropoint=25.0, return_extra=True, nthreads=0, **kw)
    got = dict(zip(NAMES, r[:7]))
    for k in ("mag", "magerr", "flux", "fluxerr", "badflag"):
        got[k] = np.asarray(got[k]).reshape(n, -1)
    _check_aper(got, dict(zip(NAMES, o[:7])), o, "big annulus, two batches", r[8]["nsky"], o[7], r[8]["mmm_iter"])
    sub = slice(0, 60)
    r = pp.aper.aper(f["image"], f["x"][sub], f["y"][sub], phpadu=f["phpadu"], zeropoint=25.0, verbose=False,
                     skyalgorithm="sigmaclipping", return_info=True, **kw)
    sky, sig, nsky = sigclip_oracle.sky(f["image"], f["x"][sub], f["y"][sub], kw["skyrad"], kw["badpix"])
    assert np.array_equal(np.asarray(r[8]["nsky"]).reshape(-1), nsky)
    assert_close(np.asarray(r[4]).reshape(-1), sky, rtol=1e-9, atol=1e-9, what="sigclip sky")
    assert_close(np.asarray(r[5]).reshape(-1), sig, rtol=1e-9, atol=1e-9, what="sigclip skyerr")


# ------------------------------------------------------------------ fourth fit variant
@pytest.mark.parametrize("radius", [3.0, 5.0])
def test_pkfit_norecenter_golden(pp, radius):
    """pkfit_norecenter.pkfit_class.pkfit_norecenter (pkfit_norecenter.py:346-657, Python-2 semantics) on the GPU against the
    reference's own outputs, scalar and batched calls, and through the unified pkfit_class."""
    from oracle import compare
    g = load_golden("golden_pkfit.npz")
    w = load_golden("golden_pkfit_norecenter.npz")["norecenter__r%d" % radius]
    cols = ("errmag", "chi", "sharp", "niter", "scale")
    want = {c: w[:, j] for j, c in enumerate(cols)}
    pk = pp.pkfit_norecenter.pkfit_class(g["image"], g["gauss"], g["psf"], float(g["ronois"]), float(g["phpadu"]))
    got = dict(zip(cols, pk.pkfit_norecenter(g["scale0"], g["x"], g["y"], g["sky"], radius)))
    rep = compare.compare_pkfit(got, want, cols, g["x"], g["y"], g["image"].shape, radius)
    assert rep["mismatches"] == 0 and rep["rank_deficient"] <= 3, rep
    i = int(np.flatnonzero(np.isfinite(w[:, 4]) & (w[:, 3] > 3))[0])
    one = pk.pkfit_norecenter(float(g["scale0"][i]), float(g["x"][i]), float(g["y"][i]), float(g["sky"][i]), radius)
    assert len(one) == 5 and one[3] == int(w[i, 3]) and abs(one[4] - w[i, 4]) <= 1e-9 * abs(w[i, 4])
    pk2 = pp.pkfit_class(g["image"], g["gauss"], g["psf"], float(g["ronois"]), float(g["phpadu"]))
    again = pk2.pkfit_norecenter(g["scale0"], g["x"], g["y"], g["sky"], radius)
    assert np.array_equal(again[3], got["niter"])
    with pytest.raises(NotImplementedError):
        pk.pkfit_fast_norecenter(1, 2, 3, 4, 5)
