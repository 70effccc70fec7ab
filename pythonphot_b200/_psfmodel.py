"""One launch evaluates every per-point piece of the DAOPHOT PSF model (csrc/pkfit_kernel.cu,
psfmodel_kernel); dao_value / daoerf / rinter slice what they need from it."""
import numpy as np

from . import _lib

ROWS = ("value", "dvdx", "dvdy", "f", "pder0", "pder1", "pder2", "pder3", "pder4", "z", "dfdx", "dfdy")


def evaluate(xx, yy, gauss, psf, device=None, table_coords=False):
    """Returns a [12, n] float64 array, rows named in ROWS."""
    torch = _lib.torch_cuda()
    dev = _lib.device_of(device)
    xx_t = _lib.to_device_f64(xx, dev)
    yy_t = _lib.to_device_f64(yy, dev, xx_t.numel())
    g_t = _lib.to_device_f64(gauss, dev)
    if g_t.numel() != 5:
        raise ValueError("gauss must have 5 elements")
    psf_h = np.ascontiguousarray(np.asarray(psf, dtype=np.float64))
    psf_t = torch.from_numpy(psf_h).to(dev)
    out_t = torch.empty((12, xx_t.numel()), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().bphot_psfmodel_device(_lib.ptr(xx_t), _lib.ptr(yy_t), xx_t.numel(), _lib.ptr(g_t), _lib.ptr(psf_t),
                                              int(psf_h.shape[1]), int(table_coords), _lib.ptr(out_t), _lib.stream_ptr(dev))
    _lib.check(rc)
    return out_t.cpu().numpy()
