#!/usr/bin/env python
"""bench.py -- stars/sec of the photometry hot path on B200 (contract: see DESIGN.md, "Measurement").

One STEP = one pass of the hot path over one synthetic exposure that is already resident in HBM:
    aper (exact pixwt weights, apr=[3,5,8], mmm sky in skyrad=[40,50])  on every star, then
    forced pkfit_norecent_noise (radius 5, noise + mask images)        on every star.
Default workload = BASELINE.json configs[3] (8192x8192 fp32 image + noise + mask, 1M stars), the largest
single-GPU configuration; the aper stage runs configs[2]'s parameters on it.  --workload aper|pkfit times
one stage alone (configs[2] / configs[3] proper).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload both|aper|pkfit]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

N > 1: one process per GPU; rank 0 builds the exposure and NCCL-broadcasts image/noise/mask once (outside the
timed steps, reported as broadcast_ms); the star list is sharded (each rank fits its own equally sized
shard: weak scaling) and the per-star result records are all-gathered INSIDE every timed step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

APR = np.array([3.0, 5.0, 8.0])
SKYRAD = (40.0, 50.0)
BADPIX = (-12000.0, 60000.0)
FIT_RADIUS = 5.0
METRIC = "stars/sec (aper+mmm sky; forced pkfit)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="both", choices=["both", "aper", "pkfit"])
    ap.add_argument("--size", type=int, default=8192, help="image is size x size")
    ap.add_argument("--stars", type=int, default=1000000, help="stars per GPU")
    ap.add_argument("--cpu-sample", type=int, default=0, help="stars in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def workload_name(a):
    return "%dx%d fp32 image+noise+u8 mask, %d stars/GPU, aper exact apr=[3,5,8] skyrad=[40,50] + forced pkfit_norecent_noise r=5" % (
        a.size, a.size, a.stars) if a.workload == "both" else (
        "%dx%d fp32, %d stars/GPU, aper exact apr=[3,5,8] skyrad=[40,50]" % (a.size, a.size, a.stars) if a.workload == "aper"
        else "%dx%d fp32 image+noise+u8 mask, %d stars/GPU, forced pkfit_norecent_noise r=5" % (a.size, a.size, a.stars))


def make_field(a, seed=1004):
    from pythonphot_b200 import synth
    return synth.make_field(a.size, a.size, a.stars, seed=seed, dtype=np.float32)


# ----------------------------------------------------------------------------- clocks
class ClockSampler(object):
    """Samples SM clock and throttle reasons of one GPU every 100 ms while the timed region runs."""

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nv = None

    def _loop(self):
        nv = self._nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80, "sync_boost": 0x10, "applications_clocks": 0x2}
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                try:
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                except Exception:
                    r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        if self._nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join(timeout=2)
        if self._nv is None:  # fall back to one nvidia-smi query
            try:
                out = subprocess.check_output(["nvidia-smi", "-i", str(self.index), "--query-gpu=clocks.sm,clocks.max.sm",
                                               "--format=csv,noheader,nounits"], timeout=10).decode().strip().split(",")
                self.samples, self.max_mhz = [int(out[0])], int(out[1])
            except Exception:
                pass
        med = int(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ----------------------------------------------------------------------------- CPU baseline (oracle port)
def cpu_baseline(field, a, nsample, threads=0):
    """Times the oracle (CPU restatement of the reference, oracle/bphot_oracle.c) with all host threads on a
    random sample of the same star list, same parameters.  Returns (stars/s, dict)."""
    from oracle import oracle
    nt = oracle.max_threads() if threads <= 0 else threads
    rng = np.random.default_rng(7)
    idx = rng.choice(field["x"].size, size=min(nsample, field["x"].size), replace=False)
    x, y, xf, yf = field["x"][idx], field["y"][idx], field["xfit"][idx], field["yfit"][idx]
    t_aper = t_fit = 0.0
    if a.workload in ("both", "aper"):
        t0 = time.perf_counter()
        oracle.aper(field["image"], x, y, phpadu=field["phpadu"], apr=APR, zeropoint=25.0, skyrad=SKYRAD, badpix=BADPIX,
                    exact=True, nthreads=nt)
        t_aper = time.perf_counter() - t0
    if a.workload in ("both", "pkfit"):
        t0 = time.perf_counter()
        oracle.pkfit(oracle.PKFIT_NORECENT, field["image"], field["gauss"], field["psf"], field["ronois"], field["phpadu"],
                     field["noise"], field["mask"], np.ones(idx.size), xf, yf, np.full(idx.size, field["sky"]), FIT_RADIUS,
                     nthreads=nt)
        t_fit = time.perf_counter() - t0
    total = t_aper + t_fit
    info = {"value": idx.size / total, "unit": "stars/s", "cores": nt, "kind": "port",
            "sample": "%d random stars of the same exposure and star list (oracle/bphot_oracle.c, OpenMP, %d threads)" % (idx.size, nt),
            "aper_stars_per_s": (idx.size / t_aper) if t_aper else None, "pkfit_stars_per_s": (idx.size / t_fit) if t_fit else None}
    return info


def run_reference(a, rank):
    """--impl reference: the reference's own CPU algorithm on the box's host cores.  The reference is pure Python and
    /root/reference does not exist on the GPU box, so the timed code is its C restatement (the oracle, pinned to the
    reference by tests/golden), with all host threads; each step is a bounded sample of the same workload."""
    if rank != 0:
        return
    from oracle import oracle
    oracle.build()
    field = make_field(a)
    nt = oracle.max_threads()
    nsample = a.cpu_sample or 400000
    for _ in range(max(a.warmup, 0)):
        cpu_baseline(field, a, max(nsample // 10, 500))
    vals, t0 = [], time.perf_counter()
    for _ in range(a.steps):
        vals.append(cpu_baseline(field, a, nsample))
    dt = time.perf_counter() - t0
    v = nsample * a.steps / sum(nsample / b["value"] for b in vals)
    base = vals[-1]
    base["value"] = v
    line = {"metric": METRIC, "value": v, "unit": "stars/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": 1e3 * dt / max(a.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "impl": "reference",
            "config": {"workload": workload_name(a), "sample_stars_per_step": nsample},
            "cpu_baseline": base, "gpu_launches": 0,
            "e2e": {"value": v, "unit": "stars/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- ours
def main():
    a = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if a.impl == "reference":
        return run_reference(a, rank)

    import torch
    import torch.distributed as dist
    import pythonphot_b200 as pp
    from pythonphot_b200 import _lib
    from pythonphot_b200.aper import aper_device

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # ---- the exposure: built on rank 0, broadcast once over NVLink (north_star), replicated in HBM
    n = a.stars
    t_b = 0.0
    if rank == 0:
        field = make_field(a)
        host = {k: torch.from_numpy(np.ascontiguousarray(field[k])).pin_memory() for k in ("image", "noise", "mask")}
    else:
        field, host = None, None
    shape = (a.size, a.size)
    img_t = host["image"].to(dev, non_blocking=True) if rank == 0 else torch.empty(shape, dtype=torch.float32, device=dev)
    noise_t = host["noise"].to(dev, non_blocking=True) if rank == 0 else torch.empty(shape, dtype=torch.float32, device=dev)
    mask_t = host["mask"].to(dev, non_blocking=True) if rank == 0 else torch.empty(shape, dtype=torch.uint8, device=dev)
    meta = [None]
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for t in (img_t, noise_t, mask_t):
            dist.broadcast(t, src=0)
        e1.record()
        torch.cuda.synchronize()
        t_b = e0.elapsed_time(e1)
        if rank == 0:
            meta = [{k: field[k] for k in ("x", "y", "xfit", "yfit", "gauss", "psf", "phpadu", "ronois", "sky")}]
        dist.broadcast_object_list(meta, src=0)
        if rank != 0:
            field = meta[0]
    # star list: `world` jittered copies of the field's list (n stars per GPU: weak scaling), known on rank 0, broadcast, then
    # sharded by image tile (pythonphot_b200/shard.py): each rank fits one contiguous Morton-ordered shard
    from pythonphot_b200 import shard
    if rank == 0:
        parts = {k: [] for k in ("x", "y", "xfit", "yfit")}
        for r in range(world):
            rng = np.random.default_rng(1000 + r)
            for k in parts:
                parts[k].append(field[k] + (rng.normal(0.0, 0.02, n) if r > 0 else 0.0))
        full = [np.concatenate(parts[k]) for k in ("x", "y", "xfit", "yfit")]
    else:
        full = [None] * 4
    cols = shard.broadcast_columns(full, world * n, dev, src=0) if world > 1 else torch.from_numpy(np.stack(full)).to(dev)
    plan = shard.ShardPlan.for_stars(cols[0].cpu().numpy(), cols[1].cpu().numpy(), a.size, a.size, world)
    assert plan.m == n
    d_x, d_y, d_xf, d_yf = (plan.take(cols[i], rank) for i in range(4))
    pin = lambda t: t.cpu().pin_memory()
    h_x, h_y, h_xf, h_yf = pin(d_x), pin(d_y), pin(d_xf), pin(d_yf)
    h_one, h_sky = torch.ones(n, dtype=torch.float64).pin_memory(), torch.full((n,), float(field["sky"]), dtype=torch.float64).pin_memory()
    d_one, d_sky = h_one.to(dev), h_sky.to(dev)
    pk = pp.pkfit_class(img_t, field["gauss"], field["psf"], field["ronois"], field["phpadu"], noise_t, mask_t)
    do_aper, do_fit = a.workload in ("both", "aper"), a.workload in ("both", "pkfit")

    bufs = {"aper": None, "fit": None}

    def step(ev=None):
        if do_aper:
            if ev:
                ev[0].record()
            bufs["aper"] = aper_device(img_t, d_x, d_y, phpadu=field["phpadu"], apr=APR, zeropoint=25.0, skyrad=SKYRAD,
                                       badpix=BADPIX, exact=True, out=bufs["aper"])
            if ev:
                ev[1].record()
        if do_fit:
            if ev:
                ev[2].record()
            bufs["fit"] = pk.fit_device(_lib.PKFIT_NORECENT, d_one, d_xf, d_yf, d_sky, FIT_RADIUS, out=bufs["fit"])
            if ev:
                ev[3].record()
        if world > 1:  # the per-star result records are packed and all-gathered, then put back in caller order (north_star)
            if ev:
                ev[4].record()
            rec = {}
            if do_aper:
                rec.update({"a_" + k: v for k, v in bufs["aper"].items()})
            if do_fit:
                rec.update({"f_" + k: bufs["fit"][k] for k in ("errmag", "chi", "sharp", "scale", "niter")})
            packed, _layout = shard.pack_records(rec)
            gathered["rec"] = {dt: plan.gather(block) for dt, block in packed.items()}
            if ev:
                ev[5].record()

    gathered = {}

    for _ in range(max(a.warmup, 3)):
        step()
    torch.cuda.synchronize()
    launches0 = _lib.launch_count()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(a.steps)]
    sampler = ClockSampler(local_rank)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_start.record()
    for k in range(a.steps):
        step(evs[k])
    t_end.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    launches = _lib.launch_count() - launches0
    ms_total = t_start.elapsed_time(t_end)
    ms_aper = float(np.mean([e[0].elapsed_time(e[1]) for e in evs])) if do_aper else 0.0
    ms_fit = float(np.mean([e[2].elapsed_time(e[3]) for e in evs])) if do_fit else 0.0
    ms_gather = float(np.mean([e[4].elapsed_time(e[5]) for e in evs])) if world > 1 else 0.0
    if world > 1:
        tt = torch.tensor([ms_total, ms_aper, ms_fit, ms_gather], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total, ms_aper, ms_fit, ms_gather = (float(v) for v in tt.tolist())
    ms_step = ms_total / a.steps
    value = world * n / (ms_step * 1e-3)
    per_rank = None
    if world > 1:  # each rank's own stage times (load balance of the shards)
        mine = torch.tensor([float(np.mean([e[0].elapsed_time(e[1]) for e in evs])) if do_aper else 0.0,
                             float(np.mean([e[2].elapsed_time(e[3]) for e in evs])) if do_fit else 0.0], dtype=torch.float64, device=dev)
        allr = torch.empty((world, 2), dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(allr, mine)
        per_rank = {"aper_ms": [round(v, 2) for v in allr[:, 0].tolist()], "pkfit_ms": [round(v, 2) for v in allr[:, 1].tolist()]}

    # ---- one extra, untimed-for-`value` pass with the library's own CUDA events around every kernel launch (on the launch
    # stream): per-kernel average launch durations for the roofline
    _lib.profile(True)
    step()
    torch.cuda.synchronize()
    ktimes = _lib.kernel_times()
    _lib.profile(False)

    # ---- end-to-end through the public API with HOST buffers (pinned), copies inside the timed region
    e2e = None
    if not a.no_e2e:
        if rank == 0:
            np_img, np_noise, np_mask = host["image"].numpy(), host["noise"].numpy(), host["mask"].numpy()
        else:  # only rank 0 holds the exposure on the host
            np_img = np_noise = np_mask = None
        nx, ny, nxf, nyf, none, nsky = (t.numpy() for t in (h_x, h_y, h_xf, h_yf, h_one, h_sky))

        def e2e_step():
            h2d = d2h = 0
            im_a, nz_a, mk_a = np_img, np_noise, np_mask
            if world > 1:
                # several GPUs: the exposure crosses PCIe once, on rank 0, and reaches the other GPUs over NVLink
                # (shard.broadcast_exposure); every rank then calls the public API on its own shard of
                # the star list with HOST star columns and gets HOST results back
                if rank == 0:
                    ex = [_lib.to_device_image(np_img, dev), _lib.to_device_image(np_noise, dev) if do_fit else None,
                          _lib.to_device_mask(np_mask, dev) if do_fit else None]
                    h2d += np_img.nbytes + (np_noise.nbytes + np_mask.nbytes if do_fit else 0)
                else:
                    ex = [torch.empty_like(img_t), torch.empty_like(noise_t) if do_fit else None,
                          torch.empty_like(mask_t) if do_fit else None]
                im_a, nz_a, mk_a = shard.broadcast_exposure(ex, src=0)
            if do_aper:
                r = pp.aper.aper(im_a, nx, ny, phpadu=field["phpadu"], apr=APR, zeropoint=25.0, skyrad=list(SKYRAD),
                                 badpix=list(BADPIX), exact=True, verbose=False)
                h2d += (np_img.nbytes if world == 1 else 0) + nx.nbytes + ny.nbytes
                d2h += sum(np.asarray(v).nbytes for v in r[:6]) + r[6].size + 8 * n  # uint8 flags + nsky/mmm_iter
            if do_fit:
                pkh = pp.pkfit_class(im_a, field["gauss"], field["psf"], field["ronois"], field["phpadu"], nz_a, mk_a)
                r = pkh.pkfit_norecent_noise(none, nxf, nyf, nsky, FIT_RADIUS)
                h2d += (np_img.nbytes + np_noise.nbytes + np_mask.nbytes if world == 1 else 0) + 4 * nxf.nbytes + field["psf"].nbytes
                d2h += sum(np.asarray(v).nbytes for v in r)
            return h2d, d2h

        e2e_step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ksteps = max(1, min(a.steps, 3))
        t0 = time.perf_counter()
        for _ in range(ksteps):
            h2d, d2h = e2e_step()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            tt = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
        e2e = {"value": world * n * ksteps / dt, "unit": "stars/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "steps": ksteps, "ms_per_step": 1e3 * dt / ksteps,
               "path": ("aper.aper + pkfit_class.pkfit_norecent_noise with numpy inputs, numpy results" if world == 1 else
                        "rank 0 uploads the exposure and NCCL-broadcasts it; every rank calls aper.aper + pkfit_class."
                        "pkfit_norecent_noise on its shard with numpy star columns, numpy results; byte counts are rank 0's")}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
        # algorithmic bytes per star (DESIGN.md "Measurement")
        nsky_mean = float(bufs["aper"]["nsky"].double().mean().item()) if do_aper else 0.0
        b_aper = 4 * (nsky_mean + 19 * 19) + 16 + 8 * (4 * APR.size + 2) + APR.size
        b_fit = 144 * (4 + 4 + 1) + 68
        stages = {}
        if do_aper:
            stages["aper"] = {"ms": ms_aper, "stars_per_s": n / (ms_aper * 1e-3), "job_stars_per_s": world * n / (ms_aper * 1e-3), "alg_bytes_per_star": b_aper,
                              "achieved_gbs": n * b_aper / (ms_aper * 1e-3) / 1e9}
        if do_fit:
            nit = bufs["fit"]["niter"]
            stages["pkfit"] = {"ms": ms_fit, "stars_per_s": n / (ms_fit * 1e-3), "job_stars_per_s": world * n / (ms_fit * 1e-3), "alg_bytes_per_star": b_fit,
                               "achieved_gbs": n * b_fit / (ms_fit * 1e-3) / 1e9,
                               "mean_niter": float(nit.double().mean().item()),
                               "frac_niter_25": float((nit == 25).double().mean().item())}
        if world > 1:
            stages["per_rank"] = per_rank
            stages["gather"] = {"ms": ms_gather, "what": "pack per dtype + NCCL all-gather + un-permute of the per-star records"}
        # per-kernel algorithmic bytes per star: the shares of SURVEY 8d's per-star figures that each kernel moves
        es = img_t.element_size()
        box = (2 * int(np.ceil(APR.max() + 0.5)) + 1) ** 2
        alg = {"aper_stage1": es * nsky_mean + 16, "aper_sums": es * box + APR.size,
               "aper_stage2": 8 * (4 * APR.size + 2) + APR.size, "pkfit": b_fit}
        try:
            traffic_db = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        except Exception:
            traffic_db = {}
        kernels = {}
        for k, (ms_k, n_k) in ktimes.items():
            if n_k == 0:
                continue
            avg = ms_k / n_k
            units = n / n_k  # stars per launch
            kernels[k] = {"launches_per_step": n_k, "avg_launch_ms": avg, "stars_per_launch": units, "alg_bytes_per_star": alg[k],
                          "achieved_gbs": units * alg[k] / (avg * 1e-3) / 1e9, "share_of_step": ms_k / ms_step}
        dom = max(kernels, key=lambda k: kernels[k]["avg_launch_ms"] * kernels[k]["launches_per_step"])
        kd = kernels[dom]
        tr = traffic_db.get(dom)
        traffic = None
        if tr:  # scale the captured per-launch DRAM bytes to this run's stars per launch
            traffic = tr["dram_bytes_per_launch"] * kd["stars_per_launch"] / tr["stars_per_launch"]
        roofline = {"bound": "hbm", "kernel": dom, "achieved": kd["achieved_gbs"], "peak": peak, "unit": "GB/s",
                    "frac": kd["achieved_gbs"] / peak, "traffic": traffic,
                    "algorithmic_bytes_per_launch": kd["stars_per_launch"] * kd["alg_bytes_per_star"],
                    "avg_launch_ms": kd["avg_launch_ms"], "peak_source": peak_src, "share_of_step": kd["share_of_step"],
                    "timing": "CUDA events recorded by libbphot on the launch stream around each launch (bphot_profile), one extra pass",
                    "traffic_source": ("profiles/" + tr["capture"]) if tr else None,
                    "note": "the path is bound by instruction issue in the sort network and by fp64 latency, not by HBM (DESIGN.md 5): "
                            "the HBM fraction is reported as the contract asks; every kernel's own numbers are in `kernels`"}
        line = {"metric": METRIC, "value": value, "unit": "stars/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "impl": "ours",
                "config": {"workload": workload_name(a), "stars_total": world * n, "l2_policy": "inputs larger than L2 (image+noise+mask = %.0f MB)" % (
                    (img_t.nbytes + noise_t.nbytes + mask_t.nbytes) / 1e6), "parallelism": "star-sharded x%d, exposure replicated" % world,
                    "broadcast_ms": t_b},
                "stages": stages, "kernels": kernels, "roofline": roofline, "clocks": clocks, "gpu_launches": int(launches), "e2e": e2e}
        if world == 1 and not a.no_cpu_baseline:
            from oracle import oracle
            oracle.build()
            line["cpu_baseline"] = cpu_baseline(field, a, a.cpu_sample or 400000)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
