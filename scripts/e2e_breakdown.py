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
