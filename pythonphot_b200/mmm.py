"""Drop-in for PythonPhot/mmm.py: the MMM sky-mode estimator, on the GPU (csrc/aper_kernel.cu, mmm_kernel)."""
import numpy as np

from . import _lib


def mmm_batch(vectors, readnoise=0, minsky=20, mxiter=50, device=None, return_iters=False):
    """mmm of several sky vectors in ONE launch (one CTA per vector).

    ``vectors``: sequence of 1-D arrays (any lengths; all float32 or they are taken as float64).
    Returns an [nvec, 3] array of (skymod, sigma, skew) (and the clipping iteration counts)."""
    torch = _lib.torch_cuda()
    dev = _lib.device_of(device)
    vecs = [np.asarray(v).reshape(-1) for v in vectors]
    nvec = len(vecs)
    if nvec == 0:
        return (np.empty((0, 3)), np.empty(0, np.int32)) if return_iters else np.empty((0, 3))
    dt = np.float32 if all(v.dtype == np.float32 for v in vecs) else np.float64
    lens = np.array([v.size for v in vecs], dtype=np.int64)
    offs = np.zeros(nvec + 1, dtype=np.int64)
    np.cumsum(lens, out=offs[1:])
    flat = np.concatenate([v.astype(dt, copy=False) for v in vecs]) if offs[-1] else np.zeros(1, dt)
    flat_t = torch.from_numpy(np.ascontiguousarray(flat)).to(dev)
    offs_t = torch.from_numpy(offs).to(dev)
    out_t = torch.empty((nvec, 3), dtype=torch.float64, device=dev)
    it_t = torch.empty(nvec, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().bphot_mmm_device(_lib.ptr(flat_t), _lib.dtype_code(flat_t), _lib.ptr(offs_t), nvec, int(lens.max()),
                                         float(readnoise or 0), int(minsky), int(mxiter), _lib.ptr(out_t), _lib.ptr(it_t),
                                         _lib.stream_ptr(dev))
    _lib.check(rc)
    out = out_t.cpu().numpy()
    return (out, it_t.cpu().numpy()) if return_iters else out


def mmm(sky_vector, highbad=False, debug=False, readnoise=False, nsky=False, integer="discrete",
        mxiter=50, minsky=20, nan=True):
    """Estimate the sky background in a stellar contaminated field (reference: mmm.py:7-314).

    Returns ``(skymod, sigma, skew)``; failures keep the reference's codes (sigma = -1.0).
    NaNs are always dropped (the reference's default ``nan=True``, mmm.py:106); ``highbad`` is not
    supported (the reference's own handling of it, mmm.py:124-125, indexes a scalar and raises)."""
    if highbad:
        raise NotImplementedError("mmm(highbad=...) is not supported (broken in the reference, mmm.py:124-125)")
    if np.iterable(minsky):
        minsky = int(np.asarray(minsky).reshape(-1)[0])
    r = mmm_batch([np.asarray(sky_vector)], readnoise=readnoise, minsky=minsky, mxiter=mxiter)[0]
    return (float(r[0]), float(r[1]), float(r[2]))
