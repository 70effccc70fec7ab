"""Golden vectors for pkfit_norecenter.pkfit_class.pkfit_norecenter (PythonPhot/pkfit_norecenter.py:346-657).

Run in the authoring container only (needs /root/reference):

    python tests/golden/gen_golden_norecenter.py

The method raises under Python 3 (``dy[good / ixx]``, pkfit_norecenter.py:435: a float index), so it is run under the
Python-2 division loader of oracle/refloader.py -- the semantics it was written for and the only ones it has.  Inputs are
those of golden_pkfit.npz (image, PSF model, star list); only the outputs are stored here.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import refloader  # noqa: E402


def main():
    g = np.load(os.path.join(HERE, "golden_pkfit.npz"))
    mod = refloader.load("pkfit_norecenter", "py2")
    out = {}
    for radius in (3.0, 5.0):
        pk = mod.pkfit_class(g["image"].astype(np.float64), g["gauss"], g["psf"], float(g["ronois"]), float(g["phpadu"]))
        rows = []
        for i in range(g["x"].size):
            with refloader.quiet(), np.errstate(all="ignore"):
                r = pk.pkfit_norecenter(float(g["scale0"][i]), float(g["x"][i]), float(g["y"][i]), float(g["sky"][i]), radius)
            rows.append([float(v) for v in r[:5]])
        out["norecenter__r%d" % radius] = np.array(rows)
    np.savez_compressed(os.path.join(HERE, "golden_pkfit_norecenter.npz"), **out)
    print("golden_pkfit_norecenter.npz", os.path.getsize(os.path.join(HERE, "golden_pkfit_norecenter.npz")))


if __name__ == "__main__":
    main()
