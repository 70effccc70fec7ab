"""find -> cntrd -> aper -> getpsf -> pkfit_noise chained through the drop-in modules (the reference notebook's flow)."""
import numpy as np
import pytest


@pytest.mark.gpu
def test_notebook_pipeline_recovers_the_planted_stars():
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("pipeline_example", os.path.join(root, "examples", "pipeline.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    r = mod.run(size=1024, nstars=1500, seed=7, verbose=False)
    f = r["field"]
    xc, yc = r["xy"]
    assert xc.size > 900  # most planted stars are detected and centroided
    # every centroid sits on a planted star (nearest neighbour within half a pixel for the bright ones)
    d2 = (xc[:, None] - f["x"][None, :]) ** 2 + (yc[:, None] - f["y"][None, :]) ** 2
    near = d2.min(axis=1)
    assert np.median(np.sqrt(near)) < 0.15
    # the PSF model recovered from the frame has the planted Gaussian widths to ~10 % (the residual table absorbs the rest)
    assert abs(r["gauss"][3] - f["gauss"][3]) < 0.2 and abs(r["gauss"][4] - f["gauss"][4]) < 0.2
    # recentring fits converge for the bulk of the stars and recover the planted fluxes of isolated bright stars
    conv = (r["niter"] > 0) & (r["niter"] < 25) & np.isfinite(r["scale"])
    assert conv.mean() > 0.8


@pytest.mark.gpu
@pytest.mark.parametrize("pinned", [False, True])
def test_aper_banded_upload_equals_resident_image(pinned):
    """A large host image goes up in row bands on a copy stream while the kernels of the first star batches already run
    (bphot_aper_device_banded); every output must equal the call on the resident image bit for bit, for page-locked and
    pageable host memory, and pkfit_class must hand the same (possibly still arriving) image to aper."""
    import torch
    import pythonphot_b200 as pp
    from pythonphot_b200 import _lib, synth
    f = synth.make_field(3000, 3100, 140000, seed=17, dtype=np.float32)
    assert f["image"].nbytes >= _lib.BAND_MIN_BYTES
    img = torch.from_numpy(f["image"]).pin_memory().numpy() if pinned else f["image"]
    kw = dict(phpadu=f["phpadu"], apr=np.array([3., 5.]), skyrad=[40, 50], badpix=[-12000, 60000], exact=True, verbose=False)
    want = pp.aper.aper(torch.from_numpy(f["image"]).cuda(), f["x"], f["y"], **kw)
    got = pp.aper.aper(img, f["x"], f["y"], **kw)
    pk = pp.pkfit_class(img, f["gauss"], f["psf"], f["ronois"], f["phpadu"], f["noise"], f["mask"])
    assert pk.image_device().events, "the engine should have uploaded the image in bands"
    via_pk = pp.aper.aper(pk.image_device(), f["x"], f["y"], **kw)
    for i in range(7):
        for other in (got, via_pk):
            assert np.array_equal(np.nan_to_num(other[i], nan=-1.0), np.nan_to_num(want[i], nan=-1.0)), "output %d differs" % i
    n = 2000
    a = pk.pkfit_norecent_noise(np.ones(n), f["xfit"][:n], f["yfit"][:n], np.full(n, f["sky"]), 5.0)
    pk2 = pp.pkfit_class(torch.from_numpy(f["image"]).cuda(), f["gauss"], f["psf"], f["ronois"], f["phpadu"],
                         torch.from_numpy(f["noise"]).cuda(), torch.from_numpy(f["mask"]).cuda())
    b = pk2.pkfit_norecent_noise(np.ones(n), f["xfit"][:n], f["yfit"][:n], np.full(n, f["sky"]), 5.0)
    for u, v in zip(a, b):
        assert np.array_equal(np.nan_to_num(u, nan=-1.0), np.nan_to_num(v, nan=-1.0))


@pytest.mark.gpu
def test_fetch_host_reads_page_locked_memory_and_refuses_pageable():
    """bphot_fetch_host: a kernel copies a page-locked host buffer (any length, any alignment of the tail) into device memory
    without using the copy engine; pageable memory is refused so that the caller falls back to cudaMemcpyAsync."""
    import ctypes as C
    import torch
    from pythonphot_b200 import _lib
    lib = _lib.lib()
    for n in (1, 17, 4096, 1000003):
        h = torch.arange(n, dtype=torch.float64).pin_memory()
        d = torch.empty(n, dtype=torch.float64, device="cuda")
        rc = lib.bphot_fetch_host(_lib.ptr(d), C.c_void_p(h.data_ptr()), n * 8, _lib.stream_ptr(d.device))
        assert rc == 0 and torch.equal(d.cpu(), h)
    hb = torch.arange(1003, dtype=torch.uint8).pin_memory()
    db = torch.zeros(1003, dtype=torch.uint8, device="cuda")
    assert lib.bphot_fetch_host(C.c_void_p(db.data_ptr() + 1), C.c_void_p(hb.data_ptr() + 1), 1001, _lib.stream_ptr(db.device)) == 0
    assert torch.equal(db.cpu()[1:1002], hb[1:1002]) and int(db[0]) == 0 and int(db[1002]) == 0
    pageable = np.arange(5000, dtype=np.float64)
    d = torch.empty(5000, dtype=torch.float64, device="cuda")
    assert lib.bphot_fetch_host(_lib.ptr(d), pageable.ctypes.data_as(C.c_void_p), 40000, _lib.stream_ptr(d.device)) == -1
    # the column helper takes either kind
    assert torch.equal(_lib.to_device_f64(pageable, d.device).cpu(), torch.from_numpy(pageable))
    pinned = torch.from_numpy(pageable).pin_memory().numpy()
    assert torch.equal(_lib.to_device_f64(pinned, d.device).cpu(), torch.from_numpy(pageable))
