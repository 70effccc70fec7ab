"""Single-GPU probe of BASELINE.json configs[4]'s pipeline at its star density: aper (exact, apr=5, skyrad=[40,50])
-> recentring pkfit_noise (radius 5) with the sky that aper returned, on a size x size tile of the 16384^2 / 10 M-star
mosaic (default 4096^2 with 625 000 stars = one GPU's sixteenth of it).  Device-resident inputs, CUDA events on the
launch stream, one JSON line.  Not the bench line (bench.py times configs[3]); this is the measurement of the
recentring variant that configs[4] names.

    python scripts/config5_probe.py [--size 4096] [--stars 625000] [--steps 3] [--warmup 3]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=4096)
    ap.add_argument("--stars", type=int, default=625000)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--unit-scale", action="store_true",
                    help="start every fit at scale=1 (README Example B's forced-fit call) instead of aper's flux / the PSF's flux")
    a = ap.parse_args()

    import torch
    import pythonphot_b200 as pp
    from pythonphot_b200 import _lib, shard, synth
    from pythonphot_b200.aper import aper_device

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    f = synth.make_field(a.size, a.size, a.stars, seed=1005, dtype=np.float32)
    img_t, noise_t, mask_t = (torch.from_numpy(np.ascontiguousarray(f[k])).to(dev) for k in ("image", "noise", "mask"))
    cols = torch.from_numpy(np.stack([f["x"], f["y"], f["xfit"], f["yfit"], f["scale"]])).to(dev)
    plan = shard.ShardPlan.for_stars(f["x"], f["y"], a.size, a.size, 1)  # image-tile order, as bench.py does
    d_x, d_y, d_xf, d_yf, d_true = (plan.take(cols[i], 0) for i in range(5))
    n = d_x.numel()
    d_one = torch.ones(n, dtype=torch.float64, device=dev)
    pk = pp.pkfit_class(img_t, f["gauss"], f["psf"], f["ronois"], f["phpadu"], noise_t, mask_t)
    bufs = {"aper": None, "fit": None}
    g = np.asarray(f["gauss"], dtype=np.float64)
    psf_flux = float(g[0] * 2.0 * np.pi * g[3] * g[4])  # integrated Gaussian part of the unit-scale PSF, ADU

    def step(ev=None):
        if ev:
            ev[0].record()
        bufs["aper"] = aper_device(img_t, d_x, d_y, phpadu=f["phpadu"], apr=np.array([5.0]), zeropoint=25.0,
                                   skyrad=(40.0, 50.0), badpix=(-12000.0, 60000.0), exact=True, out=bufs["aper"])
        if ev:
            ev[1].record()
        if a.unit_scale:
            s0 = d_one
        else:  # pkfit's `scale` argument is the star's brightness relative to the PSF's (pkfit.py docstring): seed it from aper
            fl = bufs["aper"]["flux"].reshape(n, -1)[:, 0]
            s0 = torch.where(torch.isfinite(fl) & (fl > 0), fl / psf_flux, d_one).contiguous()
        bufs["fit"] = pk.fit_device(_lib.PKFIT_NOISE, s0, d_xf, d_yf, bufs["aper"]["sky"], 5.0, out=bufs["fit"])
        if ev:
            ev[2].record()

    for _ in range(max(a.warmup, 3)):
        step()
    torch.cuda.synchronize()
    launches0 = _lib.launch_count()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(a.steps)]
    for k in range(a.steps):
        step(evs[k])
    torch.cuda.synchronize()
    ms_aper = float(np.mean([e[0].elapsed_time(e[1]) for e in evs]))
    ms_fit = float(np.mean([e[1].elapsed_time(e[2]) for e in evs]))
    ms = float(np.mean([e[0].elapsed_time(e[2]) for e in evs]))
    niter = bufs["fit"]["niter"].cpu().numpy()
    moved = torch.hypot(bufs["fit"]["x"] - d_xf, bufs["fit"]["y"] - d_yf)
    moved = moved[torch.isfinite(moved)]
    rel = (bufs["fit"]["scale"] / d_true - 1.0).abs()
    rel = rel[torch.isfinite(rel)]
    print(json.dumps({
        "workload": "%dx%d fp32 image+noise+u8 mask, %d stars (configs[4] density), aper exact apr=5 skyrad=[40,50] -> "
                    "pkfit_noise (recentring) r=5 with aper's sky" % (a.size, a.size, n),
        "steps": a.steps, "ms_per_step": ms, "stars_per_s": n / ms * 1e3,
        "aper": {"ms": ms_aper, "stars_per_s": n / ms_aper * 1e3},
        "pkfit_noise": {"ms": ms_fit, "stars_per_s": n / ms_fit * 1e3, "mean_niter": float(niter.mean()),
                        "frac_niter_25": float((niter == 25).mean()), "frac_nan": float(torch.isnan(bufs["fit"]["scale"]).float().mean()),
                        "median_shift_px": float(moved.median()) if moved.numel() else None,
                        "median_abs_rel_scale_error": float(rel.median()) if rel.numel() else None,
                        "initial_scale": "1" if a.unit_scale else "aper flux / PSF flux"},
        "gpu_launches": int(_lib.launch_count() - launches0),
    }))


if __name__ == "__main__":
    main()
