"""Reference-generated goldens at the size of BASELINE.json configs[0] and configs[1], and the reference's own speed.

Run in the authoring container only (needs /root/reference):

    python tests/golden/gen_golden_configs.py [config1] [config2] [timing]

configs[0]  README Example A (/root/reference/README.md:26-49): ``find`` -> ``aper(exact=True, apr=5, skyrad=[40,50],
            badpix=[-12000,60000])`` at the detected stars of a 2048 x 1489 frame.  The Pal5 SDSS frames are missing from
            the reference checkout, so the frame is a seeded synthetic stand-in of that shape; it is NOT
            stored (12 MB) -- the tests rebuild it from the seed and check its SHA-256 against the one stored here.
            aper is run under both readings of aper.py:246-248 (``py3`` = what the reference computes today, ``py2``).
configs[1]  README Example B (/root/reference/README.md:51-82): ``getpsf(psfrad=8, fitrad=3)`` on the brightest isolated
            stars, then ``pkfit_norecent_noise(1, x, y, sky, 5)`` on every detected star, on three frames ("g/r/i":
            three seeds) with noise and mask images.
timing      the UNMODIFIED numpy reference timed here: the README per-star loop on one core, and
            ``multiprocessing.Pool(os.cpu_count())`` over star chunks, on >= 2000-star subsamples of the bench's
            workloads.  Written to tests/golden/reference_timing.json with the host named.

Everything comes from djones1040/PythonPhot @ 7992eb6 imported in-process through oracle/refloader.py (no reference file
is edited or copied); the reference is called with fp64 arrays.
"""
import hashlib
import json
import os
import platform
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import refloader  # noqa: E402
from pythonphot_b200 import synth  # noqa: E402

C1 = dict(nrow=1489, ncol=2048, nstars=2600, seed=2001, scale_range=(4.0, 150.0), hmin=250.0, fwhm=4.1,
          roundlim=[-1.0, 1.0], sharplim=[0.2, 1.0])
C2_SEEDS = (2101, 2102, 2103)  # "g", "r", "i"
C2 = dict(nrow=1024, ncol=1024, nstars=2000, scale_range=(3.0, 150.0), hmin=250.0, fwhm=4.1)
APER_KW = dict(apr=5.0, skyrad=[40, 50], badpix=[-12000, 60000], exact=True)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def config1_field():
    """The stand-in frame of configs[0] (fp32 pixels, like an SDSS frame converted to float)."""
    f = synth.make_field(C1["nrow"], C1["ncol"], C1["nstars"], seed=C1["seed"], dtype=np.float32, scale_range=C1["scale_range"],
                         with_noise=False, edge_frac=0.01, margin=8.0)
    return f


def config2_field(seed):
    return synth.make_field(C2["nrow"], C2["ncol"], C2["nstars"], seed=seed, dtype=np.float32, scale_range=C2["scale_range"],
                            edge_frac=0.01, margin=8.0)


def gen_config1():
    f = config1_field()
    img64 = f["image"].astype(np.float64)
    out = {"image_sha256": np.array(sha(f["image"])), "phpadu": np.float64(f["phpadu"])}
    find = refloader.load("find")
    with refloader.quiet(), np.errstate(all="ignore"):
        r = find.find(img64, C1["hmin"], C1["fwhm"], C1["roundlim"], C1["sharplim"], verbose=False)
    for nm, a in zip(("x", "y", "flux", "sharp", "round"), r):
        out["find__" + nm] = np.asarray(a, dtype=np.float64)
    x, y = out["find__x"], out["find__y"]
    print("config1: find ->", x.size, "stars")
    for mode in ("py3", "py2"):
        aper = refloader.load("aper", mode)
        t0 = time.perf_counter()
        with refloader.quiet(), np.errstate(all="ignore"):
            a = aper.aper(img64, x, y, phpadu=f["phpadu"], zeropoint=25.0, verbose=False, **APER_KW)
        dt = time.perf_counter() - t0
        print("config1: aper %s %.1f s (%.2f ms/star)" % (mode, dt, 1e3 * dt / x.size))
        for nm, v in zip(("mag", "magerr", "flux", "fluxerr", "sky", "skyerr", "badflag"), a[:7]):
            out["%s__%s" % (mode, nm)] = np.asarray(v, dtype=np.float64)
        out["%s__ms_per_star" % mode] = np.float64(1e3 * dt / x.size)
    return out


def gen_config2():
    out = {}
    find = refloader.load("find")
    aper = refloader.load("aper")
    getpsf = refloader.load("getpsf")
    pk = refloader.load("pkfit_norecent_noise")
    for band, seed in zip("gri", C2_SEEDS):
        f = config2_field(seed)
        img64, nz64 = f["image"].astype(np.float64), f["noise"].astype(np.float64)
        out[band + "__image_sha256"] = np.array(sha(f["image"]))
        out[band + "__noise_sha256"] = np.array(sha(f["noise"]))
        out[band + "__mask_sha256"] = np.array(sha(f["mask"]))
        with refloader.quiet(), np.errstate(all="ignore"):
            r = find.find(img64, C2["hmin"], C2["fwhm"], [-1.0, 1.0], [0.2, 1.0], verbose=False)
            x, y = np.asarray(r[0], float), np.asarray(r[1], float)
            a = aper.aper(img64, x, y, phpadu=f["phpadu"], apr=5.0, zeropoint=25.0, skyrad=[40, 50], badpix=[-12000, 60000],
                          exact=False, verbose=False)
        mag, sky = np.asarray(a[0], float).reshape(-1), np.asarray(a[4], float).reshape(-1)
        mag = np.where(np.isfinite(mag), mag, 99.0)
        # PSF stars: the brightest detections away from the frame edge (README.md:60-66 picks them by hand)
        inner = (x > 30) & (x < C2["ncol"] - 30) & (y > 30) & (y < C2["nrow"] - 30)
        idpsf = np.argsort(np.where(inner, mag, 99.0))[:12]
        with refloader.quiet(), np.errstate(all="ignore"):
            g, p, m = getpsf.getpsf(img64, x, y, mag, sky, f["ronois"], f["phpadu"], idpsf, 8, 3, "unused.fits", verbose=False)
        g, p = np.asarray(g, np.float64), np.asarray(p, np.float64)
        print("config2 %s: %d stars, gauss %s" % (band, x.size, np.round(g, 3)))
        fit = pk.pkfit_class(img64, g, p, f["ronois"], f["phpadu"], nz64, f["mask"])
        res = np.full((x.size, 5), np.nan)
        t0 = time.perf_counter()
        with refloader.quiet(), np.errstate(all="ignore"):
            for i in range(x.size):  # README.md:73-75
                o = fit.pkfit_norecent_noise(1, x[i], y[i], sky[i], 5)
                res[i] = [float(v) for v in o[:5]]
        dt = time.perf_counter() - t0
        print("config2 %s: pkfit_norecent_noise %.1f s (%.2f ms/star)" % (band, dt, 1e3 * dt / x.size))
        out.update({band + "__x": x, band + "__y": y, band + "__mag": mag, band + "__sky": sky, band + "__idpsf": idpsf,
                    band + "__gauss": g, band + "__psf": p, band + "__psfmag": np.float64(m), band + "__fit": res,
                    band + "__ronois": np.float64(f["ronois"]), band + "__phpadu": np.float64(f["phpadu"]),
                    band + "__ms_per_star": np.float64(1e3 * dt / x.size)})
    return out


# ---------------------------------------------------------------------------- reference speed
_W = {}


def _worker_init(size, stars, seed):
    f = synth.make_field(size, size, stars, seed=seed, dtype=np.float32)
    _W["f"] = f
    _W["img"] = f["image"].astype(np.float64)
    _W["nz"] = f["noise"].astype(np.float64)
    _W["aper"] = refloader.load("aper")
    pk = refloader.load("pkfit_norecent_noise")
    _W["fit"] = pk.pkfit_class(_W["img"], f["gauss"], f["psf"], f["ronois"], f["phpadu"], _W["nz"], f["mask"])


def _worker_step(idx):
    """The bench's step on stars `idx`: aper (exact, apr=[3,5,8]) then the README forced-fit loop."""
    f = _W["f"]
    t0 = time.perf_counter()
    with refloader.quiet(), np.errstate(all="ignore"):
        _W["aper"].aper(_W["img"], f["x"][idx], f["y"][idx], phpadu=f["phpadu"], apr=np.array([3.0, 5.0, 8.0]), zeropoint=25.0,
                        skyrad=[40, 50], badpix=[-12000, 60000], exact=True, verbose=False)
        t1 = time.perf_counter()
        for i in idx:
            _W["fit"].pkfit_norecent_noise(1, f["xfit"][i], f["yfit"][i], f["sky"], 5)
    return t1 - t0, time.perf_counter() - t1


def gen_timing(size=2048, stars=62500, nsample=2048):
    """Same star density and parameters as the bench's default workload (8192^2, 1 M stars -> 2048^2, 62.5 k stars)."""
    import multiprocessing as mp
    ncpu = os.cpu_count()
    rng = np.random.default_rng(5)
    idx = rng.choice(stars, nsample, replace=False)
    _worker_init(size, stars, 1004)
    ta, tf = _worker_step(idx[: nsample // 4])
    n1 = nsample // 4
    one = {"stars": n1, "aper_s": ta, "pkfit_s": tf, "stars_per_s": n1 / (ta + tf), "aper_stars_per_s": n1 / ta,
           "pkfit_stars_per_s": n1 / tf}
    print("1 core:", one)
    chunks = np.array_split(idx, ncpu * 2)
    with mp.Pool(ncpu, initializer=_worker_init, initargs=(size, stars, 1004)) as pool:
        pool.map(_worker_step, [c[:2] for c in chunks])  # warm every worker up
        t0 = time.perf_counter()
        pool.map(_worker_step, chunks)
        dt = time.perf_counter() - t0
    allc = {"stars": nsample, "wall_s": dt, "stars_per_s": nsample / dt, "processes": ncpu}
    print("Pool(%d):" % ncpu, allc)
    return {"what": "UNMODIFIED djones1040/PythonPhot @ 7992eb6 (numpy/scipy), bench.py's default step (aper exact apr=[3,5,8] "
                    "skyrad=[40,50] + README loop of pkfit_norecent_noise r=5) on a %dx%d field of the same star density "
                    "(one star per 8x8 px), %d-star subsample" % (size, size, nsample),
            "host": {"machine": platform.machine(), "processor": platform.processor(), "cpu_count": ncpu,
                     "note": "authoring container, NOT the GPU box: /root/reference does not travel"},
            "one_core_readme_loop": one, "pool_all_cores": allc, "numpy": np.__version__}


def main(which):
    if "config1" in which:
        np.savez_compressed(os.path.join(HERE, "golden_config1.npz"), **gen_config1())
    if "config2" in which:
        np.savez_compressed(os.path.join(HERE, "golden_config2.npz"), **gen_config2())
    if "timing" in which:
        json.dump(gen_timing(), open(os.path.join(HERE, "reference_timing.json"), "w"), indent=1)
    for fn in sorted(os.listdir(HERE)):
        if fn.startswith(("golden_config", "reference_timing")):
            print(fn, os.path.getsize(os.path.join(HERE, fn)))


if __name__ == "__main__":
    main(sys.argv[1:] or ("config1", "config2", "timing"))
