"""ctypes binding of libbphot.so (C ABI declared in include/bphot.h).

The CUDA library is the only execution path of this package: if it is missing, or no CUDA device
is visible, the entry points raise -- there is no CPU fallback.
PyTorch is used for device buffers, streams and copies only.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
SO_PATH = os.path.join(CSRC, "libbphot.so")
F32, F64 = 0, 1
ERR_CAPACITY = -3
APER_NARROW_CTAS = 256  # bphot.h BPHOT_APER_NARROW_CTAS (testing aid)
PKFIT, PKFIT_NOISE, PKFIT_NORECENT, PKFIT_NORECENTER = 0, 1, 2, 3
MAX_APERTURES = 16

_lib = None

_vp, _i, _d, _ll = C.c_void_p, C.c_int, C.c_double, C.c_longlong

_SIGNATURES = {
    "bphot_version": (C.c_int, []),
    "bphot_build_flags": (C.c_char_p, []),
    "bphot_last_error": (C.c_char_p, []),
    "bphot_device_count": (C.c_int, []),
    "bphot_launch_count": (C.c_longlong, []),
    "bphot_fetch_host": (C.c_int, [_vp, _vp, _ll, _vp]),
    "bphot_pkfit_capacity": (C.c_int, [_d]),
    "bphot_profile": (C.c_int, [_i]),
    "bphot_kernel_times": (C.c_int, [_vp, _vp]),
    "bphot_aper_device": (C.c_int, [_vp, _i, _i, _i, _ll, _vp, _vp, _i, _d, _vp, _i, _d, _d, _d, _d, _d, _i, _vp, _i, _i,
                                    _d, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "bphot_aper_device_banded": (C.c_int, [_vp, _i, _i, _i, _ll, _vp, _vp, _i, _d, _vp, _i, _d, _d, _d, _d, _d, _i, _vp, _i, _i,
                                           _d, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "bphot_aper_device_noalloc": (C.c_int, [_vp, _i, _i, _i, _ll, _vp, _vp, _i, _d, _vp, _i, _d, _d, _d, _d, _d, _i, _vp, _i, _i,
                                            _d, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "bphot_aper_host": (C.c_int, [_vp, _i, _i, _i, _vp, _vp, _i, _d, _vp, _i, _d, _d, _d, _d, _d, _i, _vp, _i, _i,
                                  _d, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i]),
    "bphot_pkfit_device": (C.c_int, [_i, _vp, _i, _vp, _i, _vp, _i, _i, _ll, _ll, _ll, _vp, _vp, _i, _d, _d,
                                     _vp, _vp, _vp, _vp, _d, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "bphot_pkfit_host": (C.c_int, [_i, _vp, _i, _vp, _i, _vp, _i, _i, _vp, _vp, _i, _d, _d,
                                   _vp, _vp, _vp, _vp, _d, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i]),
    "bphot_mmm_device": (C.c_int, [_vp, _i, _vp, _i, _i, _d, _i, _i, _vp, _vp, _vp]),
    "bphot_pixwt_device": (C.c_int, [_d, _d, _d, _vp, _vp, _ll, _vp, _vp]),
    "bphot_psfmodel_device": (C.c_int, [_vp, _vp, _ll, _vp, _vp, _i, _i, _vp, _vp]),
    "bphot_find_device": (C.c_int, [_vp, _i, _i, _i, _ll, _d, _d, _d, _d, _d, _d, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "bphot_cntrd_device": (C.c_int, [_vp, _i, _i, _i, _ll, _vp, _vp, _i, _d, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "bphot_find_host": (C.c_int, [_vp, _i, _i, _i, _d, _d, _d, _d, _d, _d, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i]),
}
EXPORTS = tuple(_SIGNATURES)


class BphotError(RuntimeError):
    pass


def build(force=False, extra=""):
    """Compile libbphot.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    if force and os.path.exists(SO_PATH):
        os.remove(SO_PATH)
    cmd = ["make", "-s", "-C", CSRC, "libbphot.so"]
    if extra:
        cmd.append("EXTRA=" + extra)
    subprocess.check_call(cmd)
    return SO_PATH


def lib():
    """Load libbphot.so; raise loudly if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise BphotError("%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(or `make -C pythonphot_b200/csrc`). This package has no CPU fallback." % SO_PATH)
        handle = C.CDLL(SO_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc):
    if rc != 0:
        raise BphotError("libbphot error %d: %s" % (rc, lib().bphot_last_error().decode("utf-8", "replace")))


def build_flags():
    return lib().bphot_build_flags().decode("ascii", "replace")


def launch_count():
    return int(lib().bphot_launch_count())


KERNEL_CLASSES = ("aper_stage1", "aper_sums", "aper_stage2", "pkfit")


def profile(enable):
    """Switch per-kernel event timing in the library on/off (see include/bphot.h)."""
    return int(lib().bphot_profile(int(bool(enable))))


def kernel_times():
    """{kernel class: (summed ms, launches)} accumulated since profile(True)."""
    ms = (C.c_double * 4)()
    n = (C.c_longlong * 4)()
    check(lib().bphot_kernel_times(C.cast(ms, C.c_void_p), C.cast(n, C.c_void_p)))
    return {k: (float(ms[i]), int(n[i])) for i, k in enumerate(KERNEL_CLASSES)}


def torch_cuda():
    """Return the torch module after checking that a CUDA device is usable."""
    import torch
    if not torch.cuda.is_available():
        raise BphotError("no CUDA device is visible: pythonphot_b200 runs on the GPU only (no CPU fallback)")
    return torch


def device_of(device=None):
    torch = torch_cuda()
    if device is None:
        return torch.device("cuda", torch.cuda.current_device())
    dev = torch.device(device)
    if dev.type != "cuda":
        raise BphotError("device must be a CUDA device, got %r" % (device,))
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    return dev


_copy_streams = {}


def copy_stream(device):
    """One side stream per device for host->device uploads that should overlap kernels on the caller's stream."""
    torch = torch_cuda()
    key = str(device)
    if key not in _copy_streams:
        _copy_streams[key] = torch.cuda.Stream(device=device)
    return _copy_streams[key]


class DeviceImage(object):
    """An image in HBM that may still be arriving: ``tensor`` plus, for a banded upload, one CUDA event per band of
    ``band_rows`` rows (event k fires when rows [0, (k+1)*band_rows) are resident).  ``aper.aper`` hands the events to
    ``bphot_aper_device_banded`` so that its kernels start while the upload runs; everything else calls ``ready()``."""

    def __init__(self, tensor, events=None, band_rows=0):
        self.tensor, self.events, self.band_rows = tensor, events, band_rows

    @property
    def shape(self):
        return self.tensor.shape

    @property
    def device(self):
        return self.tensor.device

    def ready(self):
        """The tensor, after making the current stream wait for the whole upload."""
        if self.events:
            torch_cuda().cuda.current_stream(self.tensor.device).wait_event(self.events[-1])
        return self.tensor


BAND_MIN_BYTES = 32 << 20  # smaller images go up in one piece


def _upload_banded(a, device, nbands):
    """2-D contiguous host array -> DeviceImage, ``nbands`` row bands on the copy stream, one event per band."""
    torch = torch_cuda()
    host = torch.from_numpy(a)
    cur, up = torch.cuda.current_stream(device), copy_stream(device)
    t = torch.empty(a.shape, dtype=host.dtype, device=device)
    up.wait_stream(cur)
    rows = (a.shape[0] + nbands - 1) // nbands
    events = []
    with torch.cuda.stream(up):
        for r0 in range(0, a.shape[0], rows):
            t[r0:r0 + rows].copy_(host[r0:r0 + rows], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(up)
            events.append(ev)
    t.record_stream(up)
    return DeviceImage(t, events, rows)


def upload_image_banded(a, device, nbands=16, min_bytes=BAND_MIN_BYTES):
    """Host array -> ``DeviceImage`` uploaded in ``nbands`` row bands on the copy stream (asynchronous when ``a`` is
    page-locked).  Small images and CUDA tensors come back as a plain, already ordered ``DeviceImage``."""
    torch = torch_cuda()
    if isinstance(a, DeviceImage):
        return a
    if isinstance(a, torch.Tensor) or np.asarray(a).nbytes < max(min_bytes, 1):
        return DeviceImage(to_device_image(a, device))
    a = np.asarray(a)
    if a.dtype != np.float32:
        a = a.astype(np.float64, copy=False)
    a = np.ascontiguousarray(a)
    if a.ndim != 2:
        raise BphotError("image must be 2-D, got shape %s" % (a.shape,))
    return _upload_banded(a, device, nbands)


def upload_mask_banded(a, device, nbands=16, min_bytes=BAND_MIN_BYTES):
    """The same for a mask image: uint8 / bool host masks go up in bands as they are, anything else through
    ``to_device_mask`` (reduced to 0/1 on the device)."""
    torch = torch_cuda()
    if isinstance(a, DeviceImage):
        return a
    if isinstance(a, torch.Tensor) or np.asarray(a).dtype not in (np.uint8, np.bool_) or np.asarray(a).nbytes < max(min_bytes, 1):
        return DeviceImage(to_device_mask(a, device))
    a = np.ascontiguousarray(np.asarray(a).view(np.uint8))
    if a.ndim != 2:
        raise BphotError("mask must be 2-D, got shape %s" % (a.shape,))
    return _upload_banded(a, device, nbands)


def dtype_code(t):
    """BPHOT_F32/F64 for a torch tensor or numpy array of floats."""
    name = str(t.dtype)
    if name.endswith("float32"):
        return F32
    if name.endswith("float64"):
        return F64
    raise BphotError("unsupported pixel dtype %s (float32 or float64)" % name)


def to_device_image(a, device, allow=("float32", "float64")):
    """numpy array / torch tensor -> contiguous 2-D CUDA tensor (fp32 stays fp32; other dtypes -> fp64).

    A float32 image is widened pixel by pixel inside the kernels, which is bit-equivalent to the
    reference called with image.astype(float64)."""
    torch = torch_cuda()
    if isinstance(a, torch.Tensor):
        t = a
        if t.dtype not in (torch.float32, torch.float64):
            t = t.to(torch.float64)
        t = t.to(device, non_blocking=True)
    else:
        a = np.asarray(a)
        if a.dtype != np.float32:
            a = a.astype(np.float64, copy=False)
        t = torch.from_numpy(np.ascontiguousarray(a)).to(device, non_blocking=True)
    if t.dim() != 2:
        raise BphotError("image must be 2-D, got shape %s" % (tuple(t.shape),))
    return t.contiguous()


def to_device_mask(a, device):
    """mask image -> uint8 CUDA tensor, nonzero = masked (pkfit_noise.py:206-208 tests mask == 0).

    uint8 / bool masks are uploaded as they are (the kernels test ``!= 0`` themselves); other dtypes are
    reduced to 0/1 on the device after the upload, never on the host."""
    torch = torch_cuda()
    if isinstance(a, torch.Tensor):
        t = a.to(device, non_blocking=True)
        if t.dtype == torch.bool:
            t = t.view(torch.uint8)
        elif t.dtype != torch.uint8:
            t = (t != 0).view(torch.uint8)
    else:
        a = np.asarray(a)
        if a.dtype == np.bool_:
            a = a.view(np.uint8)
        t = torch.from_numpy(np.ascontiguousarray(a)).to(device, non_blocking=True)
        if t.dtype != torch.uint8:
            t = (t != 0).view(torch.uint8)
    return t.contiguous()


def to_device_f64(a, device, n=None):
    """star-list column -> contiguous fp64 CUDA vector (scalars broadcast to n)."""
    torch = torch_cuda()
    if isinstance(a, torch.Tensor):
        t = a.to(device=device, dtype=torch.float64, non_blocking=True).reshape(-1)
    else:
        h = torch.from_numpy(np.ascontiguousarray(np.atleast_1d(np.asarray(a, dtype=np.float64)).reshape(-1)))
        if h.numel() >= 4096 and h.is_pinned():
            # page-locked column: fetched by a kernel reading the host pages (bphot_fetch_host), so that it does not wait on
            # the copy engine behind an exposure upload that is still in flight
            t = torch.empty(h.shape, dtype=torch.float64, device=device)
            with torch.cuda.device(device):
                rc = lib().bphot_fetch_host(ptr(t), C.c_void_p(h.data_ptr()), h.numel() * 8, stream_ptr(device))
            if rc != 0:
                t = h.to(device, non_blocking=True)
            else:
                t._bphot_keepalive = h  # the host block must outlive the kernel; dropped with the tensor
        else:
            t = h.to(device, non_blocking=True)
    if n is not None and t.numel() != n:
        if t.numel() != 1:
            raise BphotError("star-list column has %d entries, expected %d" % (t.numel(), n))
        t = t.expand(n)
    return t.contiguous()


def to_host(tensors):
    """CUDA tensors -> numpy arrays through page-locked memory: the copies are queued back to back on the current
    stream (full PCIe rate, ~20x a pageable ``.cpu()``) and synchronised once.  The arrays own their buffers: each
    is backed by a pinned block from torch's caching host allocator, which takes the block back when the array dies."""
    torch = torch_cuda()
    hosts = []
    for t in tensors:
        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        h.copy_(t, non_blocking=True)
        hosts.append(h)
    if tensors:
        torch.cuda.current_stream(tensors[0].device).synchronize()
    return [h.numpy() for h in hosts]


def alloc_out(spec, n, device, out=None):
    """Result tensors for a launch: ``spec`` = [(name, torch dtype, tail shape)].  A caller-supplied ``out`` dict is
    checked column by column (shape, dtype, device, contiguity) before its pointers go to the C ABI: the kernels write
    n x tail elements per column without knowing how large the buffer behind the pointer is."""
    torch = torch_cuda()
    if out is None:
        return {name: torch.empty((n,) + tuple(tail), dtype=dt, device=device) for name, dt, tail in spec}
    for name, dt, tail in spec:
        t = out.get(name)
        if t is None:
            raise ValueError("out= is missing column %r" % name)
        if tuple(t.shape) != (n,) + tuple(tail) or t.dtype != dt or t.device != device or not t.is_contiguous():
            raise ValueError("out[%r] is %s %s on %s (contiguous=%s); this call needs %s %s on %s" % (
                name, tuple(t.shape), t.dtype, t.device, t.is_contiguous(), (n,) + tuple(tail), dt, device))
    return out


def host_f64(a):
    return np.ascontiguousarray(np.atleast_1d(np.asarray(a, dtype=np.float64)).reshape(-1))


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def nptr(a):
    return a.ctypes.data_as(C.c_void_p)


def stream_ptr(device):
    torch = torch_cuda()
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
