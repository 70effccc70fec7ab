"""Golden vectors for aper with sky annuli above one CTA's shared memory (the global-memory sort path of libbphot).

Run in the authoring container only (needs /root/reference):

    python tests/golden/gen_golden_bigannulus.py

The UNMODIFIED reference's aper.aper (PythonPhot/aper.py:26-433; any annulus size, aper.py:257-288 -> mmm.py:119) on a
seeded 640 x 768 frame that is NOT stored (the tests rebuild it from the seed and check its SHA-256): skyrad = [100, 180]
(about 70 000 annulus pixels per interior star) and [60, 260] (annuli cut by the frame), fp32 pixels handed over as fp64.
Also mmm.mmm (mmm.py:7) on three seeded sky vectors of 40 000 - 262 144 elements (``mmm_big``; outputs only).
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import refloader  # noqa: E402
from pythonphot_b200 import synth  # noqa: E402

SHAPE = (640, 768)
SEED = 3107
NSTARS = 40
CASES = {"r100_180": dict(apr=np.array([4.0, 9.0]), skyrad=[100, 180], badpix=[-12000, 60000], exact=True),
         "r60_260": dict(apr=np.array([4.0, 9.0]), skyrad=[60, 260], badpix=[0, 0], exact=False)}
NAMES = ("mag", "magerr", "flux", "fluxerr", "sky", "skyerr", "badflag")


def field():
    return synth.make_field(SHAPE[0], SHAPE[1], NSTARS, seed=SEED, dtype=np.float32, with_noise=False, edge_frac=0.1, margin=4.0)


def big_vectors():
    """Sky vectors longer than one CTA sorts (32 768 fp32 keys): smooth sky + a contaminated tail, some NaNs, one of integers."""
    rng = np.random.default_rng(SEED + 1)
    vecs = []
    for n, sig, integer in ((150001, 6.0, False), (40000, 2.5, True), (262144, 11.0, False)):
        v = rng.normal(300.0, sig, n)
        k = n // 12
        v[rng.integers(0, n, k)] += rng.exponential(400.0, k)
        if integer:
            v = np.round(v)
        v[rng.integers(0, n, 37)] = np.nan
        vecs.append(v)
    return vecs


def main():
    mmm = refloader.load("mmm")
    rows = []
    for v in big_vectors():
        with refloader.quiet(), np.errstate(all="ignore"):
            rows.append([float(t) for t in mmm.mmm(v.copy())[:3]])
        print("mmm n=%d ->" % v.size, rows[-1])
    f = field()
    out = {"image_sha256": np.array(hashlib.sha256(np.ascontiguousarray(f["image"]).tobytes()).hexdigest()),
           "mmm_big": np.array(rows), "x": np.asarray(f["x"], np.float64), "y": np.asarray(f["y"], np.float64), "phpadu": np.float64(f["phpadu"])}
    img64 = f["image"].astype(np.float64)
    for mode in ("py3", "py2"):
        aper = refloader.load("aper", mode)
        for cname, kw in CASES.items():
            if mode == "py2" and not kw["exact"]:
                continue
            with refloader.quiet(), np.errstate(all="ignore"):
                a = aper.aper(img64, out["x"], out["y"], phpadu=f["phpadu"], zeropoint=25.0, verbose=False, **kw)
            for nm, v in zip(NAMES, a[:7]):
                out["%s__%s__%s" % (mode, cname, nm)] = np.asarray(v, dtype=np.float64)
            print(mode, cname, "sky", np.round(np.asarray(a[4])[:4], 3), "finite fluxes", int(np.isfinite(np.asarray(a[2])).sum()))
    np.savez_compressed(os.path.join(HERE, "golden_aper_bigannulus.npz"), **out)
    print("golden_aper_bigannulus.npz", os.path.getsize(os.path.join(HERE, "golden_aper_bigannulus.npz")))


if __name__ == "__main__":
    main()
