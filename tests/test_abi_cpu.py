"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol include/bphot.h
declares; the package fails loudly without a GPU (no CPU fallback).  No compute calls."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from pythonphot_b200 import _lib
    _lib.build()
    return _lib


def test_exports_match_header(lib):
    hdr = open(os.path.join(ROOT, "include", "bphot.h")).read()
    declared = set(re.findall(r"^\s*(?:int|long long|const char\*)\s+(bphot_\w+)\s*\(", hdr, flags=re.M))
    assert declared, "no declarations parsed from include/bphot.h"
    handle = ctypes.CDLL(lib.SO_PATH)
    for name in sorted(declared):
        assert hasattr(handle, name), "libbphot.so does not export %s" % name
    assert declared == set(lib.EXPORTS), "python binding table and header disagree: %s" % (declared ^ set(lib.EXPORTS))


def test_version_and_constants(lib):
    h = lib.lib()
    hdr = open(os.path.join(ROOT, "include", "bphot.h")).read()
    assert h.bphot_version() == int(re.search(r"#define BPHOT_VERSION (\d+)", hdr).group(1))
    assert int(re.search(r"#define BPHOT_MAX_APERTURES (\d+)", hdr).group(1)) == lib.MAX_APERTURES
    assert h.bphot_launch_count() == 0


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import pythonphot_b200 as pp
    img = np.zeros((64, 64), np.float32)
    with pytest.raises(lib.BphotError):
        pp.aper.aper(img, [30.0], [30.0], verbose=False)
    with pytest.raises(lib.BphotError):
        pp.pkfit_class(img, np.ones(5), np.zeros((41, 41)), 1.0, 1.0)
    with pytest.raises(lib.BphotError):
        pp.mmm.mmm(np.arange(100.0))
    with pytest.raises(lib.BphotError):
        pp.find.find(img, 10.0, 4.0, verbose=False)
    with pytest.raises(lib.BphotError):
        pp.getpsf.getpsf(img + 1.0, np.array([32.0]), np.array([32.0]), np.array([10.0]), np.array([0.0]), 1.0, 1.0, [0], 4, 2, None,
                         verbose=False)
    # the host-buffer C entry point reports "no device" instead of computing anything
    z = np.zeros(1)
    rc = lib.lib().bphot_aper_host(lib.nptr(img), 0, 64, 64, lib.nptr(z), lib.nptr(z), 1, 1.0, lib.nptr(z), 1, 25.0, 40.0, 50.0,
                                   0.0, 0.0, 0, None, 20, 1, 0.0, 0, lib.nptr(z), lib.nptr(z), lib.nptr(z), lib.nptr(z),
                                   lib.nptr(z), lib.nptr(z), lib.nptr(z), None, None, 0)
    assert rc == -4


def test_product_package_does_not_import_oracle():
    pkg = os.path.join(ROOT, "pythonphot_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, fn)).read()
                assert "oracle" not in src.replace("# oracle", ""), "%s mentions the oracle" % fn


def test_pixel_list_capacity_bounds_every_disc(lib):
    """The fit kernels size their per-warp pixel lists from bphot_pkfit_capacity(radius); a list that is too short would
    silently drop pixels.  Brute force over 40 000 random disc centres per radius: never more pixels than the bound, and
    the bound is tight (within 12 of the observed maximum, 8 of which is rounding to a multiple of 8)."""
    h = lib.lib()
    rng = np.random.default_rng(0)
    for radius in (0.7, 1.0, 2.2, 3.0, 4.0, 5.0, 7.5, 11.0):
        cap = h.bphot_pkfit_capacity(float(radius))
        R = int(np.ceil(radius)) + 1
        g = np.arange(-R, R + 2, dtype=np.float64)
        fx, fy = rng.uniform(0, 1, 40000), rng.uniform(0, 1, 40000)
        fx[:4], fy[:4] = [0.0, 0.5, 0.0, 0.5], [0.0, 0.0, 0.5, 0.5]
        d2 = (g[None, :, None] - fx[:, None, None]) ** 2 + (g[None, None, :] - fy[:, None, None]) ** 2
        worst = int((d2 < radius * radius).sum(axis=(1, 2)).max())
        assert worst <= cap <= worst + 12, (radius, worst, cap)
    assert h.bphot_pkfit_capacity(0.0) == -1 and h.bphot_pkfit_capacity(64.0) == -1
