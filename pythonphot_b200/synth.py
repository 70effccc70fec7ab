"""Seeded synthetic star fields (numpy only) for tests, golden vectors and bench.py.

The reference ships no data (its Pal5 FITS frames are missing from the checkout,
SURVEY.md F4) and no noise/mask images, so every configuration in BASELINE.json
runs on fields made here.  A field is

    sky + sum_k scale_k * PSF(x - x_k, y - y_k) + noise

where PSF is the DAOPHOT model the reference fits (dao_value.py:95-107): a
pixel-integrated bivariate Gaussian (daoerf.py:49-61) plus a half-pixel residual
look-up table.  The table is synthetic (a smooth, zero-sum Gauss-Hermite term).
"""
import math

import numpy as np
from scipy.special import erf

SQRT2PI_REF = 2.506628275  # truncated constant used by the reference, daoerf.py:49


def make_psf(sigx=1.76, sigy=1.70, height=1000.0, psfrad=8, resid_amp=0.03):
    """Return (gauss[5], psf[npsf,npsf]) in the layout getpsf.py:142-143,292 produces."""
    n = 2 * int(psfrad + 0.5) + 1
    npsf = 2 * n + 7
    half = (npsf - 1) / 2.0
    g = (np.arange(npsf) - half) / 2.0
    xx, yy = np.meshgrid(g, g)
    psf = _resid(xx, yy, sigx, sigy, height, resid_amp)
    gauss = np.array([height, 0.0, 0.0, sigx, sigy], dtype=np.float64)
    return gauss, np.ascontiguousarray(psf, dtype=np.float64)


def _resid(dx, dy, sigx, sigy, height, amp):
    q = (dx / sigx) ** 2 + (dy / sigy) ** 2
    return height * amp * (1.0 - 0.5 * q) * np.exp(-0.5 * q) * (1.0 + 0.3 * dx / sigx * np.exp(-0.25 * q))


def _cdf(t):
    return 0.5 + 0.5 * erf(t / math.sqrt(2.0))


def star_stamps(x, y, scale, gauss, resid_amp, R):
    """Model stamps [n, 2R+1, 2R+1] and their integer origins (ix0, iy0)."""
    ix0 = np.floor(x).astype(np.int64) - R
    iy0 = np.floor(y).astype(np.int64) - R
    k = np.arange(2 * R + 1)
    dx = ix0[:, None] + k[None, :] - x[:, None]
    dy = iy0[:, None] + k[None, :] - y[:, None]
    h, sx, sy = gauss[0], gauss[3], gauss[4]
    fx = SQRT2PI_REF * sx * (_cdf((dx + 0.5) / sx) - _cdf((dx - 0.5) / sx))
    fy = SQRT2PI_REF * sy * (_cdf((dy + 0.5) / sy) - _cdf((dy - 0.5) / sy))
    st = h * fy[:, :, None] * fx[:, None, :]
    if resid_amp:
        st = st + _resid(dx[:, None, :], dy[:, :, None], sx, sy, h, resid_amp)
    st *= scale[:, None, None]
    return st, ix0, iy0


def make_field(nrow, ncol, nstars, seed, dtype=np.float32, sky=300.0, phpadu=4.0,
               ronois=5.0, margin=60.0, edge_frac=0.01, nan_frac=1e-3, sat_frac=5e-4,
               mask_frac=2e-3, scale_range=(2.0, 200.0), pos_jitter=0.05,
               with_noise=True, R=8, gauss=None, psf=None, resid_amp=0.03, chunk=20000):
    """Build one synthetic exposure.

    Returns a dict with ``image`` (dtype), ``noise`` (dtype), ``mask`` (uint8),
    true ``x``/``y``/``scale``, perturbed forced-fit positions ``xfit``/``yfit``,
    ``sky``, ``gauss``, ``psf``, ``phpadu``, ``ronois``.
    """
    rng = np.random.default_rng(seed)
    if gauss is None:
        gauss, psf = make_psf(resid_amp=resid_amp)
    n_edge = int(round(edge_frac * nstars))
    n_in = nstars - n_edge
    mx = min(margin, 0.45 * ncol)
    my = min(margin, 0.45 * nrow)
    x = rng.uniform(mx, ncol - 1 - mx, n_in)
    y = rng.uniform(my, nrow - 1 - my, n_in)
    if n_edge:
        # stars hugging / beyond the frame: exercise the clamp + badstar branches (aper.py:193-213)
        xe = rng.uniform(-3.0, ncol + 2.0, n_edge)
        ye = rng.uniform(-3.0, nrow + 2.0, n_edge)
        side = rng.integers(0, 4, n_edge)
        xe = np.where(side == 0, rng.uniform(-3.0, mx, n_edge), xe)
        xe = np.where(side == 1, rng.uniform(ncol - 1 - mx, ncol + 2.0, n_edge), xe)
        ye = np.where(side == 2, rng.uniform(-3.0, my, n_edge), ye)
        ye = np.where(side == 3, rng.uniform(nrow - 1 - my, nrow + 2.0, n_edge), ye)
        x = np.concatenate([x, xe])
        y = np.concatenate([y, ye])
    perm = rng.permutation(nstars)
    x, y = x[perm], y[perm]
    lo, hi = scale_range
    scale = np.exp(rng.uniform(math.log(lo), math.log(hi), nstars))

    model = np.full((nrow, ncol), float(sky), dtype=np.float64)
    flat = model.reshape(-1)
    k = np.arange(2 * R + 1)
    for s in range(0, nstars, chunk):
        e = min(nstars, s + chunk)
        st, ix0, iy0 = star_stamps(x[s:e], y[s:e], scale[s:e], gauss, resid_amp, R)
        cols = ix0[:, None] + k[None, :]
        rows = iy0[:, None] + k[None, :]
        ok = ((rows >= 0) & (rows < nrow))[:, :, None] & ((cols >= 0) & (cols < ncol))[:, None, :]
        idx = rows[:, :, None] * ncol + cols[:, None, :]
        np.add.at(flat, idx[ok], st[ok])

    var = np.maximum(model, 0.0) / phpadu + ronois ** 2
    out = {}
    img = np.empty((nrow, ncol), dtype=dtype)
    band = max(1, (1 << 24) // ncol)
    for r0 in range(0, nrow, band):
        r1 = min(nrow, r0 + band)
        z = rng.standard_normal((r1 - r0, ncol))
        img[r0:r1] = (model[r0:r1] + z * np.sqrt(var[r0:r1])).astype(dtype)
    if with_noise:
        out["noise"] = np.sqrt(var).astype(dtype)
        mask = np.zeros((nrow, ncol), dtype=np.uint8)
        nm = int(mask_frac * nrow * ncol)
        if nm:
            mask.reshape(-1)[rng.integers(0, nrow * ncol, nm)] = 1
        out["mask"] = mask
        nz = int(0.5 * mask_frac * nrow * ncol)
        if nz:
            out["noise"].reshape(-1)[rng.integers(0, nrow * ncol, nz)] = 0.0
    del model, var
    nn = int(nan_frac * nrow * ncol)
    if nn:
        img.reshape(-1)[rng.integers(0, nrow * ncol, nn)] = np.nan
    ns = int(sat_frac * nrow * ncol)
    if ns:
        img.reshape(-1)[rng.integers(0, nrow * ncol, ns)] = 65535.0
    out.update(image=img, x=x, y=y, scale=scale, sky=float(sky), gauss=gauss, psf=psf,
               phpadu=float(phpadu), ronois=float(ronois),
               xfit=x + rng.normal(0.0, pos_jitter, nstars),
               yfit=y + rng.normal(0.0, pos_jitter, nstars))
    return out


def make_field_device(nrow, ncol, nstars, seed, device, sky=300.0, phpadu=4.0, ronois=5.0, margin=60.0,
                      nan_frac=1e-4, sat_frac=5e-5, mask_frac=2e-4, scale_range=(0.05, 50.0), pos_jitter=0.05,
                      R=8, resid_amp=0.03, chunk=250000):
    """The same kind of exposure as ``make_field`` (sky + pixel-integrated Gaussian + residual-table stars + Poisson/read
    noise, NaN / saturated / masked pixels), built with torch ON THE DEVICE for the large configurations
    (BASELINE.json configs[4]: 16384 x 16384, 10 M stars), where the numpy generator needs minutes and ~10 GB of host
    memory.  Test / bench data only -- nothing here is on the product path.  Returns fp32 ``image`` / ``noise`` and
    uint8 ``mask`` as CUDA tensors, star columns as fp64 CUDA tensors (``x``, ``y``, ``scale``, ``xfit``, ``yfit``) and
    the PSF model (``gauss``, ``psf`` as numpy, like ``make_field``).  Deterministic for a given seed and device type."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    gauss, psf = make_psf(resid_amp=resid_amp)
    f64 = torch.float64
    u = lambda n: torch.rand(n, generator=g, device=device, dtype=f64)  # noqa: E731
    x = margin + u(nstars) * (ncol - 1 - 2 * margin)
    y = margin + u(nstars) * (nrow - 1 - 2 * margin)
    lo, hi = scale_range
    scale = torch.exp(math.log(lo) + u(nstars) * (math.log(hi) - math.log(lo)))
    model = torch.full((nrow * ncol,), float(sky), dtype=torch.float32, device=device)
    k = torch.arange(2 * R + 1, device=device)
    h, sx, sy = float(gauss[0]), float(gauss[3]), float(gauss[4])
    cdf = lambda t: 0.5 + 0.5 * torch.special.erf(t / math.sqrt(2.0))  # noqa: E731
    for s in range(0, nstars, chunk):
        e = min(nstars, s + chunk)
        xs, ys, sc = x[s:e], y[s:e], scale[s:e]
        ix0 = torch.floor(xs).long() - R
        iy0 = torch.floor(ys).long() - R
        dx = (ix0[:, None] + k[None, :]).to(f64) - xs[:, None]
        dy = (iy0[:, None] + k[None, :]).to(f64) - ys[:, None]
        fx = SQRT2PI_REF * sx * (cdf((dx + 0.5) / sx) - cdf((dx - 0.5) / sx))
        fy = SQRT2PI_REF * sy * (cdf((dy + 0.5) / sy) - cdf((dy - 0.5) / sy))
        st = h * fy[:, :, None] * fx[:, None, :]
        if resid_amp:
            ddx, ddy = dx[:, None, :], dy[:, :, None]
            q = (ddx / sx) ** 2 + (ddy / sy) ** 2
            st = st + h * resid_amp * (1.0 - 0.5 * q) * torch.exp(-0.5 * q) * (1.0 + 0.3 * ddx / sx * torch.exp(-0.25 * q))
        st = st * sc[:, None, None]
        idx = (iy0[:, None] + k[None, :])[:, :, None] * ncol + (ix0[:, None] + k[None, :])[:, None, :]
        model.index_add_(0, idx.reshape(-1), st.reshape(-1).to(torch.float32))  # stamps lie inside the frame (margin > R)
        del dx, dy, fx, fy, st, idx
    var = torch.clamp(model, min=0.0) / phpadu + ronois ** 2
    noise = torch.sqrt(var)
    image = model + torch.randn(model.shape, generator=g, device=device, dtype=torch.float32) * noise
    del model, var
    npx = nrow * ncol
    ri = lambda n: torch.randint(0, npx, (n,), generator=g, device=device)  # noqa: E731
    mask = torch.zeros(npx, dtype=torch.uint8, device=device)
    if int(mask_frac * npx):
        mask[ri(int(mask_frac * npx))] = 1
        noise[ri(int(0.5 * mask_frac * npx))] = 0.0
    if int(nan_frac * npx):
        image[ri(int(nan_frac * npx))] = float("nan")
    if int(sat_frac * npx):
        image[ri(int(sat_frac * npx))] = 65535.0
    jit = lambda n: torch.randn(n, generator=g, device=device, dtype=f64) * pos_jitter  # noqa: E731
    return dict(image=image.reshape(nrow, ncol), noise=noise.reshape(nrow, ncol), mask=mask.reshape(nrow, ncol),
                x=x, y=y, scale=scale, xfit=x + jit(nstars), yfit=y + jit(nstars), sky=float(sky), gauss=gauss, psf=psf,
                phpadu=float(phpadu), ronois=float(ronois))
