"""Where the end-to-end (host buffers in, host arrays out) time of one aper + forced-fit pass goes."""
import sys, time, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pythonphot_b200 as pp
from pythonphot_b200 import synth, _lib
from pythonphot_b200.aper import aper_device

def T(label, fn, n=3):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): r = fn()
    torch.cuda.synchronize()
    print("%-44s %8.2f ms" % (label, 1e3 * (time.perf_counter() - t0) / n), flush=True)
    return r

size, nst = int(sys.argv[1]) if len(sys.argv) > 1 else 8192, int(sys.argv[2]) if len(sys.argv) > 2 else 1000000
f = synth.make_field(size, size, nst, seed=1004, dtype=np.float32)
pin = {k: torch.from_numpy(np.ascontiguousarray(f[k])).pin_memory() for k in ("image", "noise", "mask")}
img, noise, mask = (pin[k].numpy() for k in ("image", "noise", "mask"))
dev = torch.device("cuda", 0)
print("is_pinned seen by torch:", torch.from_numpy(img).is_pinned())
img_t = T("upload image 268MB (pinned numpy)", lambda: _lib.to_device_image(img, dev))
T("upload image 268MB (pageable numpy)", lambda: _lib.to_device_image(f["image"], dev))
T("upload mask 67MB", lambda: _lib.to_device_mask(mask, dev))
x_t, y_t = _lib.to_device_f64(f["x"], dev), _lib.to_device_f64(f["y"], dev)
kw = dict(phpadu=f["phpadu"], apr=np.array([3., 5., 8.]), zeropoint=25.0, skyrad=[40, 50], badpix=[-12000, 60000], exact=True)
o = T("aper_device (kernels only)", lambda: aper_device(img_t, x_t, y_t, **kw))
T("aper results .cpu().numpy() (123 MB)", lambda: {k: v.cpu().numpy() for k, v in o.items()})
T("aper.aper() host API, device image", lambda: pp.aper.aper(img_t, f["x"], f["y"], verbose=False, **kw))
T("aper.aper() host API, host image", lambda: pp.aper.aper(img, f["x"], f["y"], verbose=False, **kw))
pk = T("pkfit_class(...) construct (3 uploads)", lambda: pp.pkfit_class(img, f["gauss"], f["psf"], f["ronois"], f["phpadu"], noise, mask))
one, sky = np.ones(nst), np.full(nst, f["sky"])
T("pk.pkfit_norecent_noise host API", lambda: pk.pkfit_norecent_noise(one, f["xfit"], f["yfit"], sky, 5.0))

# ---- round 2: the bench's end-to-end order (pkfit_class first, aper on its device image), piece by piece
print("--- round 2 path")
def e2e():
    pkh = pp.pkfit_class(img, f["gauss"], f["psf"], f["ronois"], f["phpadu"], noise, mask)
    a = pp.aper.aper(pkh.image_device(), f["x"], f["y"], verbose=False, **kw)
    b = pkh.pkfit_norecent_noise(one, f["xfit"], f["yfit"], sky, 5.0)
    return a, b
T("e2e: pkfit_class + aper(image_device) + fit", e2e)
hx = torch.from_numpy(f["x"]).pin_memory().numpy(); hy = torch.from_numpy(f["y"]).pin_memory().numpy()
T("star columns x, y -> device (pinned)", lambda: (_lib.to_device_f64(hx, dev), _lib.to_device_f64(hy, dev)))
T("star columns x, y -> device (pageable)", lambda: (_lib.to_device_f64(f["x"], dev), _lib.to_device_f64(f["y"], dev)))
o2 = dict(o); o2["badflag"] = o["badflag"].to(torch.float64)
T("_lib.to_host(aper results, 144 MB)", lambda: _lib.to_host(list(o2.values())))
big = torch.empty(256 << 20, dtype=torch.uint8, device=dev); hb = torch.empty(256 << 20, dtype=torch.uint8).pin_memory()
T("D2H 256 MiB pinned", lambda: hb.copy_(big, non_blocking=True))
T("H2D 256 MiB pinned", lambda: big.copy_(hb, non_blocking=True))
def banded_only():
    d = _lib.upload_image_banded(img, dev); d.ready(); return d
T("banded upload of the image alone", banded_only)
def aper_banded():
    d = _lib.upload_image_banded(img, dev)
    return aper_device(d.tensor, x_t, y_t, bands=(d.events, d.band_rows), **kw)
T("banded upload + aper_device (no result copy)", aper_banded)
fo = T("fit_device (kernels only)", lambda: pk.fit_device(_lib.PKFIT_NORECENT, _lib.to_device_f64(one, dev), _lib.to_device_f64(f["xfit"], dev), _lib.to_device_f64(f["yfit"], dev), _lib.to_device_f64(sky, dev), 5.0))
print("--- host-side timeline of one end-to-end step (no extra synchronisation)")
for rep_ in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    pkh = pp.pkfit_class(img, f["gauss"], f["psf"], f["ronois"], f["phpadu"], noise, mask)
    t1 = time.perf_counter()
    a = pp.aper.aper(pkh.image_device(), hx, hy, verbose=False, **kw)
    t2 = time.perf_counter()
    b = pkh.pkfit_norecent_noise(one, f["xfit"], f["yfit"], sky, 5.0)
    t3 = time.perf_counter()
    torch.cuda.synchronize()
    t4 = time.perf_counter()
    print("construct %.2f ms | aper.aper %.2f ms | fit %.2f ms | tail sync %.2f ms | total %.2f ms" % (
        1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), 1e3 * (t4 - t3), 1e3 * (t4 - t0)))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
pkh = pp.pkfit_class(img, f["gauss"], f["psf"], f["ronois"], f["phpadu"], noise, mask)
a = pp.aper.aper(pkh.image_device(), hx, hy, verbose=False, **kw)
b = pkh.pkfit_norecent_noise(one, f["xfit"], f["yfit"], sky, 5.0)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
