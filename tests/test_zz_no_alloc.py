"""bphot_aper_device_noalloc (include/bphot.h): the allocation-free aper entry point against bphot_aper_device.

No counterpart in the reference: the entry point exists for callers that may not call cudaMalloc.  Both entry points are called
through the C ABI (ctypes) on the same device tensors; integer outputs (badflag, nsky, mmm_iter) must be identical and fp64
outputs within 1e-9 relative (the bar of every parity test here) -- on two calls in a row that share the static scratch array,
the second with a different star list, so that contents left by the first call would show.  Collected last on purpose;
passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

from tests.conftest import assert_close

pytestmark = pytest.mark.gpu

APR = np.array([3.0, 5.0, 8.0])
F64_COLS = ("mag", "magerr", "flux", "fluxerr", "sky", "skyerr")
INT_COLS = ("badflag", "nsky", "mmm_iter")
SCRATCH_BYTES = 64 << 20  # BPHOT_NOALLOC_SCRATCH_BYTES


def _call(entry, img, x, y, skyrad=(40.0, 50.0), setsky=None, flags=0, exact=1):
    """One C-ABI call of `entry` (same signature for both entry points); returns (rc, dict of host arrays)."""
    import torch
    from pythonphot_b200 import _lib
    n, na, dev = x.numel(), APR.size, img.device
    out = {k: torch.full((n, na), -7.0, dtype=torch.float64, device=dev) for k in ("mag", "magerr", "flux", "fluxerr")}
    out.update({k: torch.full((n,), -7.0, dtype=torch.float64, device=dev) for k in ("sky", "skyerr")})
    out["badflag"] = torch.full((n, na), 9, dtype=torch.uint8, device=dev)
    out["nsky"] = torch.full((n,), -9, dtype=torch.int32, device=dev)
    out["mmm_iter"] = torch.full((n,), -9, dtype=torch.int32, device=dev)
    ssv = np.asarray(setsky if setsky is not None else [0.0, 0.0, 0.0], dtype=np.float64)
    with torch.cuda.device(dev):
        rc = getattr(_lib.lib(), entry)(
            _lib.ptr(img), _lib.dtype_code(img), img.shape[0], img.shape[1], img.stride(0), _lib.ptr(x), _lib.ptr(y), n,
            3.2, _lib.nptr(APR), na, 25.0, float(skyrad[0]), float(skyrad[1]), -12000.0, 60000.0,
            1 if setsky is not None else 0, _lib.nptr(ssv), 20, exact, 0.0, flags,
            _lib.ptr(out["mag"]), _lib.ptr(out["magerr"]), _lib.ptr(out["flux"]), _lib.ptr(out["fluxerr"]),
            _lib.ptr(out["sky"]), _lib.ptr(out["skyerr"]), _lib.ptr(out["badflag"]), _lib.ptr(out["nsky"]),
            _lib.ptr(out["mmm_iter"]), _lib.stream_ptr(dev))
        torch.cuda.synchronize(dev)
    return rc, {k: v.cpu().numpy() for k, v in out.items()}


def _same(got, want, what):
    for k in INT_COLS:
        assert np.array_equal(got[k], want[k]), "%s: %s differs at %s" % (what, k, np.argwhere(got[k] != want[k])[:5].tolist())
    for k in F64_COLS:
        assert_close(got[k], want[k], rtol=1e-9, what="%s %s" % (what, k))


@pytest.fixture(scope="module")
def field():
    import torch
    from pythonphot_b200 import synth
    f = synth.make_field(1024, 1024, 6000, seed=4242, dtype=np.float32)
    dev = torch.device("cuda:0")
    img = torch.from_numpy(np.ascontiguousarray(f["image"])).to(dev)
    x = torch.from_numpy(np.asarray(f["x"], dtype=np.float64)).to(dev)
    y = torch.from_numpy(np.asarray(f["y"], dtype=np.float64)).to(dev)
    return img, x, y


@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_noalloc_equals_allocating_entry_twice_in_a_row(field, dtype):
    import torch
    img, x, y = field
    img = img.to(getattr(torch, dtype))
    n1 = 3000 if dtype == "float32" else 1500  # fp64 rows are twice as long: 30 976 B per star
    # call 1: the first n1 stars; call 2 right after it on the same scratch: fewer, DIFFERENT stars
    for lo, hi in ((0, n1), (n1 + 500, n1 + 500 + n1 // 2)):
        xs, ys = x[lo:hi].contiguous(), y[lo:hi].contiguous()
        rc, got = _call("bphot_aper_device_noalloc", img, xs, ys)
        assert rc == 0, _err()
        rc, want = _call("bphot_aper_device", img, xs, ys)
        assert rc == 0, _err()
        assert (want["nsky"] > 0).any() and np.isfinite(want["flux"]).any()  # the comparison is not vacuous
        assert not (got["nsky"] == -9).any(), "an output row was never written"
        _same(got, want, "%s stars %d..%d" % (dtype, lo, hi))


def test_noalloc_other_paths(field):
    """setskyval (no sky rows at all), a small annulus (8 keys per thread) and the narrow-CTA retry pass, back to back."""
    img, x, y = field
    xs, ys = x[:2000].contiguous(), y[:2000].contiguous()
    for kw in (dict(setsky=[100.0, 3.0, 500.0]), dict(skyrad=(8.0, 14.0)), dict(skyrad=(40.0, 50.0), flags=256, exact=0)):
        rc, got = _call("bphot_aper_device_noalloc", img, xs, ys, **kw)
        assert rc == 0, _err()
        rc, want = _call("bphot_aper_device", img, xs, ys, **kw)
        assert rc == 0, _err()
        _same(got, want, repr(kw))


def test_noalloc_rejects_what_does_not_fit(field):
    img, x, y = field
    per_star = 3248 * 4 + 5248  # row of sorted keys + its index, fp32 pixels, skyrad=[40,50]
    most = (SCRATCH_BYTES - 512) // per_star
    assert most < x.numel()
    rc, out = _call("bphot_aper_device_noalloc", img, x[: most + 64].contiguous(), y[: most + 64].contiguous())
    assert rc == -3  # BPHOT_ERR_CAPACITY
    msg = _err()
    assert "static array" in msg and "at most" in msg, msg
    assert (out["nsky"] == -9).all(), "a rejected call must not touch the outputs"
    rc, _ = _call("bphot_aper_device_noalloc", img, x[:most].contiguous(), y[:most].contiguous())
    assert rc == 0, _err()


def _err():
    from pythonphot_b200 import _lib
    return _lib.lib().bphot_last_error().decode()
