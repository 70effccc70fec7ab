"""Shared implementation of the three pkfit_class variants (device-resident images + one launch per call)."""
import numpy as np

from . import _lib


class PkfitEngine(object):
    """Holds the image / noise / mask / PSF model on one CUDA device, like the reference's
    ``pkfit_class.__init__`` holds references to them (pkfit.py:121-130, pkfit_noise.py:120-129,
    pkfit_norecent_noise.py:121-130).  The upload happens once, here; every fit call afterwards moves
    only the star list."""

    def __init__(self, image, gauss, psf, ronois, phpadu, noiseim=None, maskim=None, device=None):
        torch = _lib.torch_cuda()
        if device is None and isinstance(image, _lib.DeviceImage):
            device = image.device
        if device is None and isinstance(image, torch.Tensor) and image.is_cuda:
            device = image.device
        self.device = _lib.device_of(device)
        self.f = image
        self.gauss = gauss
        self.psf = psf
        self.fnoise = noiseim
        self.fmask = maskim
        self.ronois = ronois
        self.phpadu = phpadu
        # Large host arrays go up in row bands on a copy stream of their own, image first: the constructor returns while
        # the copies run (when the arrays are page-locked), ``image_device()`` hands out the image with its band events,
        # the fits wait for all three -- so work queued in between, e.g. ``aper.aper(pk.image_device(), ...)``, overlaps the
        # noise and mask uploads.  ``DeviceImage`` arguments (e.g. from ``shard.distribute_exposure``) are taken as they are.
        self._dimg = _lib.upload_image_banded(image, self.device)
        self._dnoise = _lib.upload_image_banded(noiseim, self.device) if noiseim is not None else None
        self._dmask = _lib.upload_mask_banded(maskim, self.device) if maskim is not None else None
        self._img_t = self._dimg.tensor
        self._noise_t = self._dnoise.tensor if self._dnoise is not None else None
        self._mask_t = self._dmask.tensor if self._dmask is not None else None
        for name, t in (("noise", self._noise_t), ("mask", self._mask_t)):
            if t is not None and tuple(t.shape) != tuple(self._img_t.shape):
                raise ValueError("%s image shape %s differs from image shape %s" % (name, tuple(t.shape), tuple(self._img_t.shape)))
        self._gauss_h = _lib.host_f64(gauss)
        if self._gauss_h.size != 5:
            raise ValueError("gauss must have 5 elements (height, x0, y0, sigx, sigy)")
        psf_h = np.ascontiguousarray(np.asarray(psf.cpu() if isinstance(psf, torch.Tensor) else psf, dtype=np.float64))
        if psf_h.ndim != 2:
            raise ValueError("psf must be a 2-D look-up table")
        self._npsf = int(psf_h.shape[1])  # dao_value.py:56-57 uses shape[1]
        self._psf_t = torch.from_numpy(psf_h).to(self.device)

    # ------------------------------------------------------------------
    def image_device(self):
        """The image as it sits (or arrives) in HBM, a ``_lib.DeviceImage``, for calls that work on the same exposure --
        e.g. ``aper.aper(pk.image_device(), xc, yc, ...)`` -- so that it crosses PCIe once.  ``.ready()`` gives the
        plain CUDA tensor after the upload."""
        return self._dimg

    def _uploads_done(self):
        for d in (self._dimg, self._dnoise, self._dmask):
            if d is not None:
                d.ready()

    def fit_device(self, variant, scale_t, x_t, y_t, sky_t, radius, maxiter=25, recenter=True, out=None):
        """All star columns are fp64 CUDA vectors of equal length; returns a dict of CUDA tensors
        (errmag, chi, sharp, scale, x, y, chi2 fp64; niter int32) without synchronising."""
        torch = _lib.torch_cuda()
        n = x_t.numel()
        dev = self.device
        if variant not in (_lib.PKFIT, _lib.PKFIT_NORECENTER) and (self._noise_t is None or self._mask_t is None):
            raise ValueError("this fit needs the noise and mask images given to pkfit_class(...)")
        out = _lib.alloc_out([(k, torch.float64, ()) for k in ("errmag", "chi", "sharp", "scale", "x", "y", "chi2")] +
                             [("niter", torch.int32, ())], n, dev, out)
        img, nz, mk = self._img_t, self._noise_t, self._mask_t
        self._uploads_done()
        with torch.cuda.device(dev):
            rc = _lib.lib().bphot_pkfit_device(
                int(variant), _lib.ptr(img), _lib.dtype_code(img),
                _lib.ptr(nz), _lib.dtype_code(nz) if nz is not None else _lib.F64, _lib.ptr(mk),
                img.shape[0], img.shape[1], img.stride(0),
                nz.stride(0) if nz is not None else 0, mk.stride(0) if mk is not None else 0,
                _lib.nptr(self._gauss_h), _lib.ptr(self._psf_t), self._npsf,
                float(self.ronois), float(self.phpadu),
                _lib.ptr(scale_t), _lib.ptr(x_t), _lib.ptr(y_t), _lib.ptr(sky_t),
                float(radius), int(maxiter), int(bool(recenter)), n,
                _lib.ptr(out["errmag"]), _lib.ptr(out["chi"]), _lib.ptr(out["sharp"]), _lib.ptr(out["niter"]),
                _lib.ptr(out["scale"]), _lib.ptr(out["x"]), _lib.ptr(out["y"]), _lib.ptr(out["chi2"]),
                _lib.stream_ptr(dev))
        _lib.check(rc)
        return out

    def fit(self, variant, scale, x, y, sky, radius, maxiter=25, recenter=True, xyout=False, returnchi2=False):
        """Host-facing call: scalars or arrays in, reference-ordered tuple out
        ``(errmag, chi, sharp, niter, scale[, x, y][, chi2])``; scalars in -> scalars out."""
        torch = _lib.torch_cuda()
        scalar = not (np.iterable(x) or isinstance(x, torch.Tensor) and x.dim() > 0)
        dev = self.device
        x_t = _lib.to_device_f64(x, dev)
        n = x_t.numel()
        y_t = _lib.to_device_f64(y, dev, n)
        scale_t = _lib.to_device_f64(scale, dev, n)
        sky_t = _lib.to_device_f64(sky, dev, n)
        from .shard import FIT_SPEC, ResultArena
        arena = ResultArena(FIT_SPEC(torch), n, dev)  # the kernel writes into one byte arena: one device->host copy
        self.fit_device(variant, scale_t, x_t, y_t, sky_t, radius, maxiter=maxiter, recenter=recenter, out=arena.views)
        cols = ["errmag", "chi", "sharp", "niter", "scale"]
        if xyout:
            cols += ["x", "y"]
        if returnchi2:
            cols += ["chi2"]
        with torch.cuda.device(dev):
            h = arena.to_host()
        res = []
        for c in cols:
            a = h[c]
            if scalar:
                a = int(a[0]) if c == "niter" else float(a[0])
            res.append(a)
        return tuple(res)
