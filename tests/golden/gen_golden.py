"""Generate the committed golden vectors by running the UNMODIFIED reference.

Run in the authoring container only (needs /root/reference):

    python tests/golden/gen_golden.py

Every output in ``golden_*.npz`` comes from djones1040/PythonPhot @ 7992eb6
imported in-process through oracle/refloader.py (no reference file is edited or
copied).  Inputs (images, star lists, PSF tables) are stored beside the outputs
so the tests do not depend on the synthetic generator staying bit-stable.
The reference is called with fp64 arrays.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import refloader  # noqa: E402
from pythonphot_b200 import synth  # noqa: E402

from tests.golden.cases import APER_CASES  # noqa: E402


def gen_pixwt(rng):
    pixwt = refloader.load("pixwt")
    n = 4000
    xc = rng.uniform(3, 9, 40)
    yc = rng.uniform(3, 9, 40)
    r = rng.uniform(0.7, 8.5, 40)
    X, Y, W, P = [], [], [], []
    for i in range(40):
        x = rng.integers(-2, 14, n // 40).astype(float)
        y = rng.integers(-2, 14, n // 40).astype(float)
        if i % 4 == 1:
            y = y + rng.uniform(0, 1, y.size)  # py3-mode fractional pixel rows (aper.py:248)
        with refloader.quiet():
            w = pixwt.Pixwt(xc[i], yc[i], r[i], x, y)
        X.append(x); Y.append(y); W.append(w); P.append(np.full(x.size, i))
    # half-integer centres make x==0 inside Oneside/Arc (pixwt.py:112-125, division by zero path)
    xs = np.arange(0, 8, dtype=float)
    with refloader.quiet(), np.errstate(all="ignore"):
        w0 = pixwt.Pixwt(3.5, 3.5, 2.0, xs, xs[::-1].copy())
        doc = pixwt.Pixwt(3.23, 4.22, 3.44, np.array([5.0]), np.array([7.0]))
    return dict(xc=xc, yc=yc, r=r, pid=np.concatenate(P), x=np.concatenate(X), y=np.concatenate(Y),
                w=np.concatenate(W), half_x=xs, half_y=xs[::-1].copy(), half_w=w0, doc=doc)


def gen_mmm(rng):
    mmm = refloader.load("mmm")
    vecs, outs = [], []

    def add(v, **kw):
        with refloader.quiet():
            o = mmm.mmm(np.asarray(v, dtype=np.float64), **kw)
        vecs.append(np.asarray(v, dtype=np.float64))
        outs.append([float(o[0]), float(o[1]), float(o[2]), float(kw.get("readnoise", 0) or 0),
                     float(kw.get("minsky", 20))])

    for n in (25, 60, 400, 2827, 2840, 5000):
        v = rng.normal(300, 9, n)
        add(v)
        v2 = v.copy()
        k = max(1, n // 12)
        v2[rng.integers(0, n, k)] += rng.exponential(400, k)  # star contamination
        add(v2)
        v3 = np.round(v2)  # integer-valued data: ties in the sorted vector
        add(v3)
    add(rng.normal(300, 9, 15))  # too few (mmm.py:109-112)
    v = rng.normal(300, 9, 500); v[::7] = np.nan
    add(v)  # NaNs dropped (mmm.py:106)
    add(np.full(100, 7.0))  # zero variance
    add(np.concatenate([np.full(30, 1.0), np.full(30, 1000.0)]))  # bimodal
    add(rng.normal(300, 9, 800), readnoise=6.0)
    add(rng.normal(300, 9, 90), minsky=80)
    add(np.concatenate([rng.normal(300, 2, 40), rng.normal(5000, 2, 45)]), minsky=20)
    lens = np.array([v.size for v in vecs])
    return dict(vec=np.concatenate(vecs), lens=lens, out=np.array(outs))


def gen_psfmodel(rng, gauss, psf):
    daoerf = refloader.load("daoerf")
    rinter = refloader.load("rinter")
    dao_value = refloader.load("dao_value")
    n = 500
    xx = rng.uniform(-6.5, 6.5, n)
    yy = rng.uniform(-6.5, 6.5, n)
    g2 = np.array([gauss[0], 0.13, -0.21, gauss[3], gauss[4]])
    with refloader.quiet():
        f, pder = daoerf.daoerf(xx, yy, g2)
        half = (psf.shape[0] - 1) / 2.0
        z, dfdx, dfdy = rinter.rinter(psf, 2 * xx + half, 2 * yy + half, deriv=True)
        v, dvx, dvy = dao_value.dao_value(xx, yy, g2, psf, deriv=True)
        # edge guard (dao_value.py:70-84): one point off the table zeroes everything
        xe = xx[:10].copy(); xe[3] = 30.0
        ve, dvxe, dvye = dao_value.dao_value(xe, yy[:10], g2, psf, deriv=True)
    return dict(xx=xx, yy=yy, gauss=g2, psf=psf, f=f, pder=pder, z=z, dfdx=dfdx, dfdy=dfdy,
                v=v, dvx=dvx, dvy=dvy, xe=xe, ve=np.asarray(ve, float) * np.ones(10))


def gen_aper(field):
    img64 = field["image"].astype(np.float64)
    x, y = field["x"], field["y"]
    out = dict(image=field["image"], x=x, y=y, phpadu=np.float64(field["phpadu"]))
    names = ("mag", "magerr", "flux", "fluxerr", "sky", "skyerr", "badflag")
    for mode in ("py3", "py2"):
        aper = refloader.load("aper", mode)
        for cname, kw in APER_CASES:
            if mode == "py2" and not kw.get("exact"):
                continue  # identical to py3
            kw = dict(kw)
            if "minsky" in kw:  # aper.py:148 calls len(minsky): an int raises, a 1-element array works
                kw["minsky"] = np.array([kw["minsky"]])
            with refloader.quiet(), np.errstate(all="ignore"):
                r = aper.aper(img64, x, y, phpadu=field["phpadu"], zeropoint=25.0, verbose=False, **kw)
            for nm, a in zip(names, r[:7]):
                a = np.asarray(a, dtype=np.float64)
                if nm in ("sky", "skyerr"):
                    a = a.reshape(-1)
                else:
                    a = a.reshape(x.size, -1)
                out["%s__%s__%s" % (mode, cname, nm)] = a
    return out


def _rows(ref):
    """Ragged reference tuples -> [n, width] (some early returns drop chi2 / x,y; pad with NaN)."""
    w = max(len(r) for r in ref)
    out = np.full((len(ref), w), np.nan)
    for i, r in enumerate(ref):
        out[i, :len(r)] = [float(np.ravel(v)[0]) for v in r]
    return out


def gen_pkfit(field, n=160):
    img64 = field["image"].astype(np.float64)
    nz64 = field["noise"].astype(np.float64)
    mask = field["mask"]
    gauss, psf = field["gauss"], field["psf"]
    x, y = field["xfit"][:n].copy(), field["yfit"][:n].copy()
    x[5] = np.nan  # NaN position branch (pkfit_norecent_noise.py:156-164, pkfit_noise.py:162-170)
    scale0 = field["scale"][:n] * np.where(np.arange(n) % 3 == 0, 1.0, 0.6)
    sky = np.full(n, field["sky"])
    out = dict(image=field["image"], noise=field["noise"], mask=mask, gauss=gauss, psf=psf, x=x, y=y,
               scale0=scale0, sky=sky, ronois=np.float64(field["ronois"]), phpadu=np.float64(field["phpadu"]))
    for radius in (3.0, 5.0):
        tag = "r%d" % radius
        m = refloader.load("pkfit_norecent_noise")
        pk = m.pkfit_class(img64, gauss, psf, field["ronois"], field["phpadu"], nz64, mask)
        with refloader.quiet(), np.errstate(all="ignore"):
            ref = [pk.pkfit_norecent_noise(scale0[i], x[i], y[i], sky[i], radius, returnchi2=True) for i in range(n)]
        out["norecent__%s" % tag] = _rows(ref)  # errmag,chi,sharp,niter,scale,chi2
        m = refloader.load("pkfit_noise")
        pk = m.pkfit_class(img64, gauss, psf, field["ronois"], field["phpadu"], nz64, mask)
        with refloader.quiet(), np.errstate(all="ignore"):
            ref = [pk.pkfit_noise(scale0[i], x[i], y[i], sky[i], radius, xyout=True) for i in range(n)]
        out["noise__%s" % tag] = _rows(ref)  # errmag,chi,sharp,niter,scale,x,y
        m = refloader.load("pkfit")
        pk = m.pkfit_class(img64, gauss, psf, field["ronois"], field["phpadu"])
        for rec in (True, False):
            with refloader.quiet(), np.errstate(all="ignore"):
                ref = [pk.pkfit(scale0[i], x[i], y[i], sky[i], radius, xyout=True, recenter=rec) for i in range(n)]
            out["pkfit_%s__%s" % ("rec" if rec else "norec", tag)] = np.array(
                [[float(np.ravel(v)[0]) for v in r] for r in ref])
    return out


def main():
    rng = np.random.default_rng(20261104)
    np.savez_compressed(os.path.join(HERE, "golden_pixwt.npz"), **gen_pixwt(rng))
    np.savez_compressed(os.path.join(HERE, "golden_mmm.npz"), **gen_mmm(rng))
    field = synth.make_field(230, 250, 140, seed=11, dtype=np.float32, margin=52.0, edge_frac=0.12,
                             nan_frac=4e-4, sat_frac=3e-4)
    np.savez_compressed(os.path.join(HERE, "golden_psfmodel.npz"), **gen_psfmodel(rng, field["gauss"], field["psf"]))
    np.savez_compressed(os.path.join(HERE, "golden_aper.npz"), **gen_aper(field))
    field2 = synth.make_field(200, 220, 160, seed=12, dtype=np.float32, margin=12.0, edge_frac=0.08,
                              nan_frac=2e-4, sat_frac=2e-4)
    np.savez_compressed(os.path.join(HERE, "golden_pkfit.npz"), **gen_pkfit(field2))
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
