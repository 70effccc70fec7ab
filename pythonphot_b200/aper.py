"""Drop-in for PythonPhot/aper.py: concentric aperture photometry with mmm sky, on the GPU.

``aper(image, xc, yc, ...)`` keeps the reference signature (aper.py:26-29) and returns the same
8-tuple ``(mag, magerr, flux, fluxerr, sky, skyerr, badflag, outstr)`` (aper.py:433); the per-star
loop (aper.py:217-417) runs as one CUDA launch (csrc/aper_kernel.cu) through libbphot's C ABI.
"""
import ctypes as C

import numpy as np

from . import _lib


class OutStr(object):
    """The reference's per-star summary strings (aper.py:407-409), formatted on access so that a
    million-star call does not pay for a million ``%`` operations nobody reads."""

    def __init__(self, xc, yc, mag, magerr):
        self._xc, self._yc, self._mag, self._magerr = xc, yc, mag, magerr

    def __len__(self):
        return len(self._xc)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        return '%.3f,%.3f :' % (self._xc[i], self._yc[i]) + \
            '  '.join(['%.4f+-%.4f' % (m, e) for m, e in zip(self._mag[i], self._magerr[i])])

    def __iter__(self):
        return (self[i] for i in range(len(self)))


def _minsky_value(minsky):
    # aper.py:148: the default [] means 20
    if minsky is None:
        return 20
    if np.iterable(minsky):
        return 20 if len(minsky) == 0 else int(np.asarray(minsky).reshape(-1)[0])
    return int(minsky)


def aper_device(image_t, xc_t, yc_t, phpadu=1, apr=5, zeropoint=25, skyrad=(40, 50), badpix=(0, 0),
                setskyval=None, minsky=20, exact=False, readnoise=0, compat='py3', out=None, skyalgorithm='mmm',
                _narrow_ctas=False, bands=None):
    """Device-resident core: ``image_t`` [nrow, ncol] and ``xc_t``/``yc_t`` [nstars] are CUDA tensors.

    Enqueues one kernel on the current stream and returns a dict of CUDA tensors
    (mag, magerr, flux, fluxerr [nstars, naper] fp64; sky, skyerr [nstars] fp64; badflag uint8
    [nstars, naper]; nsky, mmm_iter int32 [nstars]) without synchronising.  ``out`` may carry a
    dict from a previous call with the same shapes to reuse its buffers (checked: shape, dtype, device).
    ``bands`` = ``(events, band_rows)`` of an image that is still being uploaded band by band (``_lib.DeviceImage``):
    the kernels of each star batch wait only for the band they read (``bphot_aper_device_banded``)."""
    torch = _lib.torch_cuda()
    if compat not in ('py3', 'py2'):
        raise ValueError("compat must be 'py3' or 'py2'")
    dev = image_t.device
    apr_h = _lib.host_f64(apr)
    naper, n = apr_h.size, xc_t.numel()
    if naper < 1 or naper > _lib.MAX_APERTURES:
        raise ValueError("between 1 and %d apertures are supported, got %d" % (_lib.MAX_APERTURES, naper))
    assert yc_t.numel() == n, 'xc and yc arrays must be identical length.'
    if setskyval is not None:  # aper.py:156-159
        ssv = _lib.host_f64(setskyval)
        if ssv.size == 1:
            ssv = np.array([ssv[0], 0., 1.])
        assert ssv.size == 3, 'Keyword SETSKYVAL must contain 1 or 3 elements'
        use_setsky = 1
    else:
        ssv, use_setsky = np.zeros(3), 0
    skyrad = np.asarray(skyrad, dtype=np.float64).reshape(-1)
    badpix = np.asarray(badpix, dtype=np.float64).reshape(-1)
    f64, na = torch.float64, (naper,)
    out = _lib.alloc_out([('mag', f64, na), ('magerr', f64, na), ('flux', f64, na), ('fluxerr', f64, na), ('sky', f64, ()),
                          ('skyerr', f64, ()), ('badflag', torch.uint8, na), ('nsky', torch.int32, ()),
                          ('mmm_iter', torch.int32, ())], n, dev, out)
    ev_arr, nbands, band_rows = None, 0, 0
    if bands is not None and bands[0]:
        nbands, band_rows = len(bands[0]), int(bands[1])
        ev_arr = (C.c_void_p * nbands)(*[C.c_void_p(e.cuda_event) for e in bands[0]])
    with torch.cuda.device(dev):
        rc = _lib.lib().bphot_aper_device_banded(
            _lib.ptr(image_t), _lib.dtype_code(image_t), image_t.shape[0], image_t.shape[1], image_t.stride(0),
            _lib.ptr(xc_t), _lib.ptr(yc_t), n, float(phpadu), _lib.nptr(apr_h), naper, float(zeropoint),
            float(skyrad[0]), float(skyrad[1]), float(badpix[0]), float(badpix[1]), use_setsky, _lib.nptr(ssv),
            int(minsky), int(bool(exact)), float(readnoise or 0),
            (1 if compat == 'py2' else 0) | (2 if str(skyalgorithm).startswith('sigmaclip') else 0)
            | (_lib.APER_NARROW_CTAS if _narrow_ctas else 0),
            _lib.ptr(out['mag']), _lib.ptr(out['magerr']), _lib.ptr(out['flux']), _lib.ptr(out['fluxerr']),
            _lib.ptr(out['sky']), _lib.ptr(out['skyerr']), _lib.ptr(out['badflag']), _lib.ptr(out['nsky']),
            _lib.ptr(out['mmm_iter']), C.cast(ev_arr, C.c_void_p) if ev_arr is not None else C.c_void_p(0), nbands, band_rows,
            _lib.stream_ptr(dev))
    _lib.check(rc)
    return out


def aper(image, xc, yc, phpadu=1, apr=5, zeropoint=25,
         skyrad=[40, 50], badpix=[0, 0], setskyval=None, minsky=[],
         skyalgorithm='mmm', exact=False, readnoise=0,
         verbose=True, debug=False, compat='py3', device=None, return_info=False):
    """Compute concentric aperture photometry on one or more stars (reference: aper.py:26-433).

    Same positional/keyword arguments and the same return tuple as the reference (``skyalgorithm`` = 'mmm' or
    'sigmaclipping', aper.py:275-288).  Extras
    (keyword only in spirit): ``compat`` selects how aper.py:246-248 is read -- 'py3' = as the
    reference runs today (pixel y-centre = row + col/nx when exact=True), 'py2' = the Python-2
    integer row it was written for; ``device`` picks the CUDA device; ``return_info`` appends a dict
    with ``nsky`` and ``mmm_iter`` per star.  ``image`` may be a numpy array, a CUDA tensor or the ``DeviceImage`` a
    ``pkfit_class`` made of the same exposure (``pk.image_device()``: the image then crosses PCIe once for both); a
    float32 image is widened per pixel on the GPU, i.e. results equal the reference called with
    ``image.astype(float64)``.  ``apr`` may also be a list (the reference needs scalar/ndarray).
    """
    if skyalgorithm != 'mmm' and not str(skyalgorithm).startswith('sigmaclip'):  # aper.py:275: anything else means mmm
        skyalgorithm = 'mmm'
    torch = _lib.torch_cuda()
    on_dev = isinstance(image, _lib.DeviceImage) or (isinstance(image, torch.Tensor) and image.is_cuda)
    dev = _lib.device_of(device if device is not None else (image.device if on_dev else None))
    # aper.py:134-148
    if not np.iterable(xc):
        xc, yc = np.array([xc]), np.array([yc])
    assert len(xc) == len(yc), 'xc and yc arrays must be identical length.'
    nstars = len(xc)
    # a large host image goes up in row bands on a copy stream while the kernels already work on the rows that arrived
    dimg = _lib.upload_image_banded(image, dev)
    img_t = dimg.tensor
    xc_t = _lib.to_device_f64(xc, dev)
    yc_t = _lib.to_device_f64(yc, dev)
    # the kernels write into one byte arena, which comes back in ONE device->host copy
    from .shard import APER_SPEC, ResultArena
    naper = _lib.host_f64(apr).size
    arena = ResultArena(APER_SPEC(torch, naper) + [('badflag_f', torch.float64, (naper,))], nstars, dev)
    o = aper_device(img_t, xc_t, yc_t, phpadu=phpadu, apr=apr, zeropoint=zeropoint, skyrad=skyrad, badpix=badpix,
                    setskyval=setskyval, minsky=_minsky_value(minsky), exact=exact, readnoise=readnoise,
                    compat=compat, skyalgorithm=skyalgorithm, bands=(dimg.events, dimg.band_rows) if dimg.events else None,
                    out={k: v for k, v in arena.views.items() if k != 'badflag_f'})
    arena.views['badflag_f'].copy_(o['badflag'])  # the reference stores 0./1. (aper.py:174); cast on the device, not on the host
    with torch.cuda.device(dev):
        h = arena.to_host()
    mag, magerr, flux, fluxerr = h['mag'], h['magerr'], h['flux'], h['fluxerr']
    sky, skyerr = h['sky'], h['skyerr']
    badflag = h['badflag_f']
    xh = np.asarray(xc.cpu() if isinstance(xc, torch.Tensor) else xc, dtype=np.float64).reshape(-1)
    yh = np.asarray(yc.cpu() if isinstance(yc, torch.Tensor) else yc, dtype=np.float64).reshape(-1)
    outstr = OutStr(xh, yh, mag, magerr)
    if verbose:
        nbad = int(((xh < 0.5) | (xh > img_t.shape[1] - 1.5) | (yh < 0.5) | (yh > img_t.shape[0] - 1.5)).sum())
        if nbad:  # aper.py:211-213
            print('WARNING [aper.py] - ' + str(nbad) + ' star positions outside image')
    if nstars == 1:  # aper.py:419-427
        sky, skyerr = sky[0], skyerr[0]
        mag, magerr, flux, fluxerr, badflag = mag[0], magerr[0], flux[0], fluxerr[0], badflag[0]
        outstr = outstr[0]
    res = (mag, magerr, flux, fluxerr, sky, skyerr, badflag, outstr)
    if return_info:
        res = res + ({'nsky': h['nsky'], 'mmm_iter': h['mmm_iter']},)
    return res
