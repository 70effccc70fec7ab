"""Parity comparison of the CUDA path's outputs with the oracle's (or the reference goldens').

TEST INFRASTRUCTURE ONLY (like everything under oracle/): used by tests/test_gpu_parity.py and by bench.py's in-run
parity check (the `parity` key of the bench line).  The bar is north_star's: counts, flags and iteration numbers
bit-exact; fp64 values within 1e-9 relative (aperture fluxes additionally 1e-7 ADU absolute: they are differences
of 1e4..1e6-ADU sums).  Two classes of stars are excluded from VALUE comparison, counted and bounded:
fragile apertures (oracle flag) and rank-deficient fits (fewer than 4 in-frame pixels inside the fitting radius).
"""
import numpy as np

APER_NAMES = ("mag", "magerr", "flux", "fluxerr", "sky", "skyerr", "badflag")


def close_mask(a, b, rtol=1e-9, atol=0.0):
    """Elementwise verdict: NaN pattern equal, infinities equal, finite values within atol + rtol*|b|."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    na, nb = np.isnan(a), np.isnan(b)
    ia, ib = np.isinf(a), np.isinf(b)
    ok = (na == nb) & (ia == ib)
    fin = ~(na | nb | ia | ib)
    with np.errstate(invalid="ignore"):
        ok &= np.where(fin, np.abs(a - b) <= atol + rtol * np.abs(b), True)
        ok &= np.where(ia & ib, a == b, True)
    return ok


def underdetermined(x, y, radius, shape):
    """Stars with fewer than 4 in-frame pixels inside the fitting radius: their 3x3 normal matrix is rank
    deficient, and whether LAPACK meets an exactly zero pivot (-> NaN return, pkfit.py:383-406) or a
    1e-17 one (-> garbage step) hinges on the last bit of erf(); neither implementation is 'right'."""
    ny_, nx_ = shape
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    R = int(np.ceil(radius)) + 1
    xi, yi = np.floor(np.nan_to_num(x)).astype(np.int64), np.floor(np.nan_to_num(y)).astype(np.int64)
    cnt = np.zeros(x.size, dtype=np.int64)
    for oy in range(-R, R + 1):
        for ox in range(-R, R + 1):
            px, py = xi + ox, yi + oy
            inside = (px >= 0) & (px < nx_) & (py >= 0) & (py < ny_)
            cnt += inside & (((px - x) ** 2 + (py - y) ** 2) < radius ** 2)
    return (cnt < 4) & np.isfinite(x) & np.isfinite(y)


def _aper_values_ok(got, cand, sel):
    """[n, naper] bool: the four value columns of `got` agree with `cand` (dict of [n, naper]) where `sel`."""
    ok = np.ones(sel.shape, dtype=bool)
    for nm in ("flux", "fluxerr", "mag", "magerr"):
        atol = 1e-7 if nm in ("flux", "fluxerr") else 1e-9
        ok &= close_mask(got[nm], cand[nm], 1e-9, atol)
    return ok | ~sel


def compare_aper(got, want, fragile, nsky=None, want_nsky=None, mmm_iter=None, want_iter=None, alt=None):
    """got / want: dicts with APER_NAMES ([n, naper] and [n] arrays).  `fragile` [n, naper] bool from the oracle.

    Robust apertures: badflag exact, values within tolerance.  Fragile apertures (a NaN / out-of-range pixel sits where
    Pixwt returns 0 +- 1e-16, so the reference's own result depends on the sign of a rounding error): with `alt` (the
    oracle's tenth output) the GPU result must equal ONE of the outcomes the reference's arithmetic admits --
    the oracle's own; every fragile pixel left out; a finite out-of-range one let in (same values, flag 1); a non-finite
    one let in (pixel sum zeroed, aper.py:366, flag 1) -- without `alt` they are only counted.
    Returns {"stars", "mismatches", "fragile", "fragile_frac", "fragile_checked", "first_bad": {column: [indices]}}."""
    n = np.asarray(want["sky"]).size
    shp = (n, -1)
    fragile = np.asarray(fragile, dtype=bool).reshape(shp)
    g = {k: np.asarray(got[k], dtype=np.float64).reshape(shp) for k in ("flux", "fluxerr", "mag", "magerr", "badflag")}
    w = {k: np.asarray(want[k], dtype=np.float64).reshape(shp) for k in ("flux", "fluxerr", "mag", "magerr", "badflag")}
    gb, wb = g["badflag"] != 0, w["badflag"] != 0
    bad = {}
    flag_mism = gb != wb
    robust_flag_mism = flag_mism & ~fragile
    if robust_flag_mism.any():
        bad["badflag"] = np.nonzero(robust_flag_mism.any(axis=1))[0]
    ok_ap = ~flag_mism & ~fragile  # apertures whose value is comparable one to one
    vals_ok = _aper_values_ok(g, w, ok_ap)
    if not vals_ok.all():
        bad["values"] = np.nonzero(~vals_ok.all(axis=1))[0]
    checked = 0
    if alt is not None and fragile.any():
        kind = np.asarray(alt["kind"]).reshape(shp)
        a5 = {k: np.asarray(alt[k], dtype=np.float64).reshape(shp) for k in ("flux", "fluxerr", "mag", "magerr", "badflag")}
        # outcome "a non-finite fragile pixel got in": the weighted sum is zeroed, the rest of aper.py:383-403 as usual
        apr = np.asarray(alt["apr"], dtype=np.float64).reshape(1, -1)
        area = np.pi * apr * apr
        sky = np.asarray(want["sky"], dtype=np.float64).reshape(-1, 1)
        skysig = np.asarray(want["skyerr"], dtype=np.float64).reshape(-1, 1)
        if alt.get("setsky") is not None:
            nsk = np.full((n, 1), float(alt["setsky"][2]))
        else:
            nsk = np.maximum(np.asarray(alt["nsky"], dtype=np.float64).reshape(-1, 1), 1.0)
        with np.errstate(invalid="ignore", divide="ignore"):
            skyvar = skysig * skysig
            zf = 0.0 - sky * area
            ze = np.sqrt(area * skyvar + np.abs(zf) / alt["phpadu"] + (skyvar / nsk) * (area * area))
            zm = np.where(zf > 0, alt["zeropoint"] - 2.5 * np.log10(np.where(zf > 0, zf, 1.0)), np.nan)
            zme = np.where(zf > 0, 1.0857 * ze / np.where(zf > 0, zf, 1.0), np.nan)
        zero = {"flux": zf + 0 * area, "fluxerr": ze, "mag": zm, "magerr": zme}
        own = _aper_values_ok(g, w, fragile) & (gb == wb)
        excl = _aper_values_ok(g, a5, fragile) & ((gb == (a5["badflag"] != 0)) | (gb & ((kind & 4) != 0)))
        zeroed = _aper_values_ok(g, zero, fragile) & gb & ((kind & 2) != 0)
        fr_ok = (own | excl | zeroed) | ~fragile
        checked = int(fragile.sum())
        if not fr_ok.all():
            bad["fragile"] = np.nonzero(~fr_ok.all(axis=1))[0]
    for nm in ("sky", "skyerr"):
        c = close_mask(np.asarray(got[nm]).reshape(-1), np.asarray(want[nm]).reshape(-1), 1e-9, 1e-9)
        if not c.all():
            bad[nm] = np.nonzero(~c)[0]
    for nm, g_, w_ in (("nsky", nsky, want_nsky), ("mmm_iter", mmm_iter, want_iter)):
        if g_ is not None and w_ is not None:
            c = np.asarray(g_).reshape(-1).astype(np.int64) == np.asarray(w_).reshape(-1).astype(np.int64)
            if not c.all():
                bad[nm] = np.nonzero(~c)[0]
    stars_bad = np.unique(np.concatenate([v for v in bad.values()])) if bad else np.zeros(0, np.int64)
    return {"stars": int(n), "mismatches": int(stars_bad.size), "fragile": int(fragile.sum()),
            "fragile_frac": float(fragile.mean()) if fragile.size else 0.0, "fragile_checked": checked,
            "fragile_flag_differs": int((flag_mism & fragile).sum()),
            "first_bad": {k: [int(i) for i in v[:8]] for k, v in bad.items()}}


def compare_pkfit(got, want, cols, x, y, shape, radius):
    """got / want: dicts of [n] arrays; `cols` = the columns to compare (niter exact, the rest 1e-9 relative; a NaN in
    want's x / y / chi2 means the reference dropped that column on an early return).  Rank-deficient stars are skipped
    and counted.  Off-frame stars see only the far Gaussian tail, where erf implementations differ at ~1e-9: their
    `sharp` gets 1e-6."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    n = x.size
    skip = underdetermined(x, y, radius, shape)
    keep0 = ~skip
    bad = {}
    c = (np.asarray(got["niter"]).astype(np.int64) == np.asarray(want["niter"]).astype(np.int64)) | skip
    if not c.all():
        bad["niter"] = np.nonzero(~c)[0]
    ny_, nx_ = shape
    off = ~((x >= 0) & (x <= nx_ - 1) & (y >= 0) & (y <= ny_ - 1))
    for nm in cols:
        if nm == "niter":
            continue
        a, w = np.asarray(got[nm], dtype=np.float64), np.asarray(want[nm], dtype=np.float64)
        keep = keep0.copy()
        if nm in ("x", "y", "chi2"):
            keep &= ~np.isnan(w)
        c = close_mask(a, w, 1e-9, 1e-12)
        if nm == "sharp":
            c = np.where(off, close_mask(a, w, 1e-6, 1e-12), c)
        c = c | ~keep
        if not c.all():
            bad[nm] = np.nonzero(~c)[0]
    stars_bad = np.unique(np.concatenate([v for v in bad.values()])) if bad else np.zeros(0, np.int64)
    return {"stars": int(n), "mismatches": int(stars_bad.size), "rank_deficient": int(skip.sum()),
            "rank_deficient_frac": float(skip.mean()) if n else 0.0,
            "first_bad": {k: [int(i) for i in v[:8]] for k, v in bad.items()}}
