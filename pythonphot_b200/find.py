"""Drop-in for PythonPhot/find.py (DAOFIND): detect stars, return centroids and shape statistics.

The convolution, the local-maximum search and the per-candidate statistics (find.py:188-346) run as
CUDA kernels (csrc/find_kernel.cu) behind ``bphot_find_device``; this module only normalises arguments,
sizes the output columns and prints what the reference prints."""
import ctypes as C

import numpy as np

from . import _lib


def find_device(image_t, hmin, fwhm, roundlim=(-1.0, 1.0), sharplim=(0.2, 1.0), max_out=None):
    """``image_t``: 2-D CUDA tensor (fp32 / fp64).  Returns ``(cols, counts)``: ``cols`` a [5, nstars] fp64
    CUDA tensor (x, y, flux, sharp, round) and ``counts`` = dict(nstars, ngood, badsharp, badround, badcntrd)."""
    torch = _lib.torch_cuda()
    dev = image_t.device
    cap = int(max_out) if max_out else max(4096, min(image_t.numel() // 64, 1 << 22))
    counts = (C.c_int * 5)()
    while True:
        out = torch.empty((5, cap), dtype=torch.float64, device=dev)
        with torch.cuda.device(dev):
            rc = _lib.lib().bphot_find_device(
                _lib.ptr(image_t), _lib.dtype_code(image_t), image_t.shape[0], image_t.shape[1], image_t.stride(0),
                float(hmin), float(fwhm), float(roundlim[0]), float(roundlim[1]), float(sharplim[0]), float(sharplim[1]),
                cap, _lib.ptr(out[0]), _lib.ptr(out[1]), _lib.ptr(out[2]), _lib.ptr(out[3]), _lib.ptr(out[4]),
                C.cast(counts, C.c_void_p), _lib.stream_ptr(dev))
        if rc == _lib.ERR_CAPACITY and counts[0] > cap:
            cap = int(counts[0])  # the library reports the size needed; run again with columns that hold it
            continue
        _lib.check(rc)
        break
    names = ("nstars", "ngood", "badsharp", "badround", "badcntrd")
    cnt = {k: int(counts[i]) for i, k in enumerate(names)}
    return out[:, :cnt["nstars"]], cnt


def find(image,
         hmin, fwhm,
         roundlim=[-1.0, 1.0],
         sharplim=[0.2, 1.0],
         doprint=False,
         verbose=True, device=None):
    """Find positive brightness perturbations (stars) in an image (reference: find.py:8-394).

    Same arguments and return as the reference: ``(x, y, flux, sharp, roundness)`` in the reference's
    star order, or ``None`` when nothing passes.  ``image`` may be a numpy array or a CUDA tensor; integer
    images are widened to fp64 like the reference does (find.py:92), fp32 is widened per pixel on the GPU.
    ``doprint`` writes the same table to 'find.prt' (or the given name)."""
    torch = _lib.torch_cuda()
    dev = _lib.device_of(device if device is not None else (image.device if isinstance(image, torch.Tensor) and image.is_cuda else None))
    if len(np.shape(image)) != 2:
        print('ERROR - Image array (first parameter) must be 2 dimensional')
    img_t = _lib.to_device_image(image, dev)
    n_y, n_x = img_t.shape
    if verbose:
        print('Input Image Size is ' + str(n_x) + ' by ' + str(n_y))
    if fwhm < 0.5:
        print('ERROR - Supplied FWHM must be at least 0.5 pixels')
    if verbose:
        print('Beginning convolution of image')
    cols, cnt = find_device(img_t, hmin, fwhm, roundlim, sharplim)
    if verbose:
        print('Finished convolution of image')
    if cnt["ngood"] == 0:
        return None  # find.py:213-217, 236-263: no local maxima above the threshold
    if verbose:
        print(str(cnt["ngood"]) + ' local maxima located above threshold')
        print(' No. of sources rejected by SHARPNESS criteria', cnt["badsharp"])
        print(' No. of sources rejected by ROUNDNESS criteria', cnt["badround"])
        print(' No. of sources rejected by CENTROID  criteria', cnt["badcntrd"])
    if cnt["nstars"] == 0:
        return None  # find.py:373
    with torch.cuda.device(dev):
        h = _lib.to_host([cols.contiguous()])[0]
    x, y, flux, sharp, roundness = (np.ascontiguousarray(h[i]) for i in range(5))
    if doprint:
        import time
        fname = 'find.prt' if doprint == 1 or doprint is True else doprint
        with open(fname, 'w') as fout:
            print(' Program: FIND ' + time.asctime(time.localtime(time.time())), file=fout)
            print(' Threshold above background:', hmin, file=fout)
            print(' Approximate FWHM:', fwhm, file=fout)
            print(' Sharpness Limits: Low', sharplim[0], '  High', sharplim[1], file=fout)
            print(' Roundness Limits: Low', roundlim[0], '  High', roundlim[1], file=fout)
            print(' No of sources above threshold', cnt["ngood"], file=fout)
            print(' No. of sources rejected by SHARPNESS criteria', cnt["badsharp"], file=fout)
            print(' No. of sources rejected by ROUNDNESS criteria', cnt["badround"], file=fout)
            print(' No. of sources rejected by CENTROID  criteria', cnt["badcntrd"], file=fout)
            print('     STAR       X       Y     FLUX     SHARP    ROUND', file=fout)
            for i in range(x.size):
                print(i + 1, x[i], y[i], flux[i], sharp[i], roundness[i], file=fout)
    return (x, y, flux, sharp, roundness)
