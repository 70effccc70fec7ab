"""Drop-in for PythonPhot/pixwt.py: exact circle / unit-pixel overlap area, on the GPU."""
import numpy as np

from . import _lib


def Pixwt(xc, yc, r, x, y, device=None):
    """Fraction of the unit pixel centred at (x, y) inside the circle (xc, yc, r) (pixwt.py:8-48).

    ``x``, ``y`` scalars or equal-shaped arrays; the result has their shape (scalar in -> float out;
    the reference's scalar call raises AttributeError at pixwt.py:62, its array call is what aper uses)."""
    torch = _lib.torch_cuda()
    dev = _lib.device_of(device)
    scalar = not np.iterable(x)
    xa = np.atleast_1d(np.asarray(x, dtype=np.float64))
    ya = np.atleast_1d(np.asarray(y, dtype=np.float64))
    xa, ya = np.broadcast_arrays(xa, ya)
    shape = xa.shape
    x_t = _lib.to_device_f64(np.ascontiguousarray(xa), dev)
    y_t = _lib.to_device_f64(np.ascontiguousarray(ya), dev)
    out_t = torch.empty_like(x_t)
    with torch.cuda.device(dev):
        rc = _lib.lib().bphot_pixwt_device(float(xc), float(yc), float(r), _lib.ptr(x_t), _lib.ptr(y_t), x_t.numel(),
                                           _lib.ptr(out_t), _lib.stream_ptr(dev))
    _lib.check(rc)
    out = out_t.cpu().numpy().reshape(shape)
    return float(out[0]) if scalar else out


def Intarea(xc, yc, r, x0, x1, y0, y1):
    """Area of overlap of the circle (xc, yc, r) and the UNIT rectangle [x0,x1]x[y0,y1] (pixwt.py:215-235);
    only unit pixels are supported (x1-x0 = y1-y0 = 1), which is the only way the reference calls it."""
    x0, x1, y0, y1 = (np.asarray(v, dtype=np.float64) for v in (x0, x1, y0, y1))
    if not (np.allclose(x1 - x0, 1.0) and np.allclose(y1 - y0, 1.0)):
        raise NotImplementedError("Intarea: only unit pixels are supported on the GPU path")
    return Pixwt(xc, yc, r, 0.5 * (x0 + x1), 0.5 * (y0 + y1))
