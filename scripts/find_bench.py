"""Device-side timing of find (DAOFIND) and cntrd on a resident frame: per-call ms, algorithmic GB/s, kernel split."""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pythonphot_b200 import synth
from pythonphot_b200.find import find_device
from pythonphot_b200.cntrd import cntrd_device

size, nst = int(sys.argv[1]) if len(sys.argv) > 1 else 8192, int(sys.argv[2]) if len(sys.argv) > 2 else 1000000
f = synth.make_field(size, size, nst, seed=1004, dtype=np.float32, with_noise=False)
img = torch.from_numpy(f["image"]).cuda()
peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))).get("hbm_gbs", 6553.0) if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6553.0
for fwhm in (4.1, 9.0):
    cols, cnt = find_device(img, 60.0, fwhm)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        cols, cnt = find_device(img, 60.0, fwhm)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    npix = size * size
    alg = npix * (4 + 8 + 8 + 1 + 1)  # image read, h write, h read, flag write, flag read (select)
    print(json.dumps({"op": "find", "fwhm": fwhm, "frame": "%dx%d fp32" % (size, size), "ms": ms, "stars": cnt["nstars"], "local_maxima": cnt["ngood"],
                      "Mpix_per_s": npix / ms / 1e3, "alg_GBps": alg / ms / 1e6, "frac_of_hbm_peak": alg / ms / 1e6 / peak}))
x, y = cols[0].contiguous(), cols[1].contiguous()
cntrd_device(img, x, y, 4.1); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    cntrd_device(img, x, y, 4.1)
e1.record(); torch.cuda.synchronize()
print(json.dumps({"op": "cntrd", "stars": int(x.numel()), "ms": e0.elapsed_time(e1) / 5, "Mstars_per_s": x.numel() / (e0.elapsed_time(e1) / 5) / 1e3}))
