"""Drop-in for PythonPhot/cntrd.py: centroids by derivative search (the early DAOPHOT FIND algorithm)."""
import numpy as np

from . import _lib

_MESSAGES = {1: 'Position %s too near edge of image', 2: 'Position %s moved too near edge of image',
             3: 'Unable to compute X centroid around position %s', 4: 'Computed X centroid for position %s out of range',
             5: 'Unable to compute Y centroid around position %s', 6: 'Computed Y centroid for position %s out of range',
             7: 'No finite pixel around position %s'}


def cntrd_device(image_t, x_t, y_t, fwhm, extendbox=0, keepcenter=False):
    """CUDA tensors in, ``(xcen, ycen, status)`` CUDA tensors out (status: BPHOT_CNTRD_* of include/bphot.h)."""
    torch = _lib.torch_cuda()
    dev = image_t.device
    if hasattr(extendbox, '__len__'):
        ex, ey = extendbox
    else:
        ex = ey = extendbox
    n = x_t.numel()
    xcen = torch.empty(n, dtype=torch.float64, device=dev)
    ycen = torch.empty(n, dtype=torch.float64, device=dev)
    status = torch.empty(n, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().bphot_cntrd_device(_lib.ptr(image_t), _lib.dtype_code(image_t), image_t.shape[0], image_t.shape[1],
                                           image_t.stride(0), _lib.ptr(x_t), _lib.ptr(y_t), n, float(fwhm), int(ex or 0), int(ey or 0),
                                           int(bool(keepcenter)), _lib.ptr(xcen), _lib.ptr(ycen), _lib.ptr(status), _lib.stream_ptr(dev))
    _lib.check(rc)
    return xcen, ycen, status


def cntrd(img, x, y,
          fwhm, verbose=True,
          debug=False,
          extendbox=False,
          keepcenter=False, device=None):
    """Compute the centroid of stars using a derivative search (reference: cntrd.py:7-231).

    Same arguments and return ``(xcen, ycen)``; positions the reference gives up on come back as -1, -1.
    One launch for the whole list (csrc/find_kernel.cu, cntrd_kernel)."""
    torch = _lib.torch_cuda()
    dev = _lib.device_of(device if device is not None else (img.device if isinstance(img, torch.Tensor) and img.is_cuda else None))
    img_t = _lib.to_device_image(img, dev)
    x_t = _lib.to_device_f64(x, dev)
    y_t = _lib.to_device_f64(y, dev, x_t.numel())
    with torch.cuda.device(dev):
        xc, yc, st = _lib.to_host(list(cntrd_device(img_t, x_t, y_t, fwhm, extendbox or 0, keepcenter)))
    if verbose and st.any():
        xs, ys = np.atleast_1d(np.asarray(x, dtype=np.float64)), np.atleast_1d(np.asarray(y, dtype=np.float64))
        for i in np.flatnonzero(st):
            print(_MESSAGES[int(st[i])] % (str(xs[i]) + ' ' + str(ys[i])))
    if xc.size == 1:  # cntrd.py:229-230: one position in, two scalars out
        return float(xc[0]), float(yc[0])
    return xc, yc
