#!/usr/bin/env python
"""bench.py -- stars/sec of the photometry hot path on B200 (contract: see DESIGN.md, "Measurement").

One STEP = one pass of the hot path over one synthetic exposure that is already resident in HBM:
    aper (exact pixwt weights, apr=[3,5,8], mmm sky in skyrad=[40,50])  on every star, then
    forced pkfit_norecent_noise (radius 5, noise + mask images)        on every star.
Default workload = BASELINE.json configs[3] (8192x8192 fp32 image + noise + mask, 1M stars), the largest
single-GPU configuration; the aper stage runs configs[2]'s parameters on it.  --workload aper|pkfit times
one stage alone (configs[2] / configs[3] proper); --workload config5 runs BASELINE.json configs[4] alone
(16384x16384 mosaic, 10M stars in total, aper apr=5 -> recentring pkfit_noise, star-sharded); at N > 1 the default
line carries the same run as its `config5` block.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload both|aper|pkfit|config5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

N > 1: one process per GPU; rank 0 builds the exposure and NCCL-broadcasts image/noise/mask once (outside the
timed steps, reported as broadcast_ms); the star list is sharded (each rank fits its own equally sized
shard: weak scaling) and the per-star result records are all-gathered INSIDE every timed step.

Every line proves what it timed: `parity` compares a sample of the timed run's own results with the CPU oracle
(counts / flags / iteration numbers exact, values 1e-9), and at N > 1 `shard_check` recomputes a sample on one GPU
and requires the gathered records to be bit-identical.  A mismatch makes the run exit non-zero.
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
REF_BUDGET_S = 240.0  # the whole --impl reference run (warm-up + steps) is sized to end within this


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="both", choices=["both", "aper", "pkfit", "config5"])
    ap.add_argument("--size", type=int, default=8192, help="image is size x size")
    ap.add_argument("--stars", type=int, default=1000000, help="stars per GPU")
    ap.add_argument("--cpu-sample", type=int, default=0, help="stars in the CPU baseline / parity sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-config5", action="store_true", help="N > 1: skip the config5 block")
    ap.add_argument("--image-dtype", default="float32", choices=["float32", "float64"],
                    help="pixel type of image and noise (float64 = the reference's native dtype; the default workload is fp32)")
    ap.add_argument("--c5-size", type=int, default=16384)
    ap.add_argument("--c5-stars", type=int, default=10000000, help="config5: stars in TOTAL (strong scaling)")
    return ap.parse_args()


def host_threads():
    """Host threads this process may use.  NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1 to every worker."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def workload_name(a):
    name = "%dx%d fp32 image+noise+u8 mask, %d stars/GPU, aper exact apr=[3,5,8] skyrad=[40,50] + forced pkfit_norecent_noise r=5" % (
        a.size, a.size, a.stars) if a.workload == "both" else (
        "%dx%d fp32, %d stars/GPU, aper exact apr=[3,5,8] skyrad=[40,50]" % (a.size, a.size, a.stars) if a.workload == "aper"
        else "%dx%d fp32 image+noise+u8 mask, %d stars/GPU, forced pkfit_norecent_noise r=5" % (a.size, a.size, a.stars))
    return name.replace("fp32", "fp64") if a.image_dtype == "float64" else name


def config_dict(a, world):
    """The same dict from both arms (the driver compares them): it names the workload, nothing measured goes in here."""
    es = 8 if a.image_dtype == "float64" else 4
    mb = a.size * a.size * (es + es + 1) / 1e6
    return {"workload": workload_name(a), "stars_total": world * a.stars,
            "l2_policy": "inputs larger than L2 (image+noise+mask = %.0f MB)" % mb,
            "parallelism": "star-sharded x%d, exposure replicated" % world}


def make_field(a, seed=1004):
    from pythonphot_b200 import synth
    return synth.make_field(a.size, a.size, a.stars, seed=seed, dtype=np.dtype(a.image_dtype).type)


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


# ----------------------------------------------------------------------------- CPU oracle legs
def oracle_pass(field, workload, x, y, xf, yf, nt, variant=None, apr=APR, fit_sky=None, fit_scale=None):
    """One pass of the oracle (CPU restatement of the reference, oracle/bphot_oracle.c, OpenMP over `nt` threads) over the
    given stars of the exposure, same parameters as the GPU step.  Returns (seconds aper, seconds fit, outputs)."""
    from oracle import oracle
    t_aper = t_fit = 0.0
    out = {}
    if workload in ("both", "aper", "config5"):
        t0 = time.perf_counter()
        out["aper"] = oracle.aper(field["image"], x, y, phpadu=field["phpadu"], apr=apr, zeropoint=25.0, skyrad=SKYRAD,
                                  badpix=BADPIX, exact=True, nthreads=nt, return_extra=True)
        t_aper = time.perf_counter() - t0
    if workload in ("both", "pkfit", "config5"):
        t0 = time.perf_counter()
        v = oracle.PKFIT_NORECENT if variant is None else variant
        sky = np.full(xf.size, field["sky"]) if fit_sky is None else fit_sky
        scale = np.ones(xf.size) if fit_scale is None else fit_scale
        out["fit"] = oracle.pkfit(v, field["image"], field["gauss"], field["psf"], field["ronois"], field["phpadu"],
                                  field["noise"], field["mask"], scale, xf, yf, sky, FIT_RADIUS, nthreads=nt)
        t_fit = time.perf_counter() - t0
    return t_aper, t_fit, out


def cpu_baseline(field, a, nsample, threads=0, keep=False):
    """Times the oracle with all host threads on a random sample of the same star list, same parameters.
    Returns the `cpu_baseline` dict (and, with keep=True, the sample indices and the oracle's outputs for the parity check)."""
    nt = host_threads() if threads <= 0 else threads
    rng = np.random.default_rng(7)
    idx = rng.choice(field["x"].size, size=min(nsample, field["x"].size), replace=False)
    t_aper, t_fit, out = oracle_pass(field, a.workload, field["x"][idx], field["y"][idx], field["xfit"][idx], field["yfit"][idx], nt)
    total = t_aper + t_fit
    numpy_ref = None
    try:  # the UNMODIFIED numpy reference on this workload, timed where /root/reference exists: a constant
        t = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_timing.json")))
        numpy_ref = {"one_core_stars_per_s": t["one_core_readme_loop"]["stars_per_s"], "pool_stars_per_s": t["pool_all_cores"]["stars_per_s"],
                     "pool_processes": t["pool_all_cores"]["processes"], "host": "authoring container (%d vCPU), not this box" % t["host"]["cpu_count"],
                     "source": "tests/golden/reference_timing.json"}
    except Exception:
        pass
    info = {"value": idx.size / total, "unit": "stars/s", "cores": nt, "kind": "port", "numpy_reference": numpy_ref,
            "sample": "%d random stars of the same exposure and star list (oracle/bphot_oracle.c, OpenMP, %d threads)" % (idx.size, nt),
            "aper_stars_per_s": (idx.size / t_aper) if t_aper else None, "pkfit_stars_per_s": (idx.size / t_fit) if t_fit else None}
    return (info, idx, out) if keep else info


def parity_report(got_aper, got_fit, want, xs, ys, xfs, yfs, shape, fit_cols):
    """`parity` key: the GPU step's own results for the sampled stars against the oracle's (oracle/compare.py)."""
    from oracle import compare
    rep = {"stars": int(xs.size), "mismatches": 0, "tolerance": "counts/flags/niter exact; values 1e-9 rel (+1e-7 ADU on fluxes)"}
    if got_aper is not None:
        o = want["aper"]
        w = dict(zip(compare.APER_NAMES, o[:7]))
        r = compare.compare_aper(got_aper, w, o[9], got_aper["nsky"], o[7], got_aper["mmm_iter"], o[8], alt=o[10])
        rep["aper"] = r
        rep["mismatches"] += r["mismatches"]
        rep["fragile"] = r["fragile"]
    if got_fit is not None:
        r = compare.compare_pkfit(got_fit, want["fit"], fit_cols, xfs, yfs, shape, FIT_RADIUS)
        rep["pkfit"] = r
        rep["mismatches"] += r["mismatches"]
        rep["rank_deficient"] = r["rank_deficient"]
    return rep


def run_reference(a, rank, world):
    """--impl reference: the reference's own CPU algorithm on the box's host cores.  The reference is pure Python and
    /root/reference does not exist on the GPU box, so the timed code is its C restatement (the oracle, pinned to the
    reference by tests/golden), with all host threads (from the affinity mask: torchrun's OMP_NUM_THREADS=1 is ignored);
    each step is a bounded sample of the same workload, sized from a probe so that the run ends within REF_BUDGET_S."""
    if rank != 0:
        return
    from oracle import oracle
    oracle.build()
    field = make_field(a)
    nt = host_threads()
    probe = cpu_baseline(field, a, 2000, nt)  # also warms the thread pool up
    budget = REF_BUDGET_S / (max(a.steps, 1) + 0.1 * max(a.warmup, 0))
    nsample = a.cpu_sample or int(min(400000, max(2000, 0.8 * probe["value"] * budget)))
    nsample = min(nsample, a.stars)
    for _ in range(max(a.warmup, 0)):
        cpu_baseline(field, a, max(nsample // 10, 500), nt)
    vals, t0 = [], time.perf_counter()
    for _ in range(a.steps):
        vals.append(cpu_baseline(field, a, nsample, nt))
    dt = time.perf_counter() - t0
    v = nsample * a.steps / sum(nsample / b["value"] for b in vals)
    base = vals[-1]
    base["value"] = v
    line = {"metric": METRIC, "value": v, "unit": "stars/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": 1e3 * dt / max(a.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "impl": "reference", "config": config_dict(a, max(world, a.gpus)),
            "sample_stars_per_step": nsample, "cpu_baseline": base, "gpu_launches": 0,
            "e2e": {"value": v, "unit": "stars/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- helpers of the product arm
def bitwise_equal(t1, t2):
    import torch
    a, b = t1.contiguous(), t2.contiguous()
    return a.shape == b.shape and a.dtype == b.dtype and bool(torch.equal(a.view(torch.uint8), b.view(torch.uint8)))


def to_np(d, idx_t=None):
    return {k: (v if idx_t is None else v.index_select(0, idx_t)).cpu().numpy() for k, v in d.items()}


def timed(fn, steps, world, dist, torch, dev):
    """barrier + synchronize, `steps` calls of fn(k), synchronize + barrier; returns total ms (max over ranks)."""
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(steps):
        fn(k)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        tt = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
    return ms


# ----------------------------------------------------------------------------- config 5
def run_config5(a, rank, world, dev, steps, warmup):
    """BASELINE.json configs[4]: 16384 x 16384 fp32 mosaic + noise + u8 mask, 10 M stars in TOTAL (strong scaling),
    aper(exact, apr=5, skyrad=[40,50]) then the recentring pkfit_noise (radius 5; sky and starting scale from aper),
    star-sharded through ShardedPhotometry.  Returns the `config5` dict (rank 0) or None."""
    import torch
    import torch.distributed as dist
    from pythonphot_b200 import _lib, shard, synth
    size, ntot = a.c5_size, a.c5_stars
    t0 = time.perf_counter()
    if rank == 0:
        # star brightness is scaled down with the density (one star per %.0f px^2): scale log-uniform in [0.01, 10] x a
        # 1000-ADU PSF over a 300-ADU sky, 1e-4 NaN, 5e-5 saturated, 2e-4 masked pixels
        f = synth.make_field_device(size, size, ntot, seed=1005, device=dev, scale_range=(0.01, 10.0))
        torch.cuda.synchronize()
    else:
        f = None
    t_gen = time.perf_counter() - t0
    meta = shard.broadcast_header({k: f[k] for k in ("gauss", "psf", "phpadu", "ronois", "sky")} if rank == 0 else None)
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    sp = shard.ShardedPhotometry(f["image"] if rank == 0 else None, meta["gauss"], meta["psf"], meta["ronois"], meta["phpadu"],
                                 f["noise"] if rank == 0 else None, f["mask"] if rank == 0 else None, device=dev)
    e1.record()
    torch.cuda.synchronize()
    t_bcast = e0.elapsed_time(e1)
    cols = shard.broadcast_columns([f[k] for k in ("x", "y")] if rank == 0 else [None, None], ntot, dev, src=0)
    x_t, y_t = cols[0], cols[1]
    plan = sp.plan(x_t, y_t)
    apr5 = np.array([5.0])
    kw = dict(phpadu=meta["phpadu"], apr=apr5, zeropoint=25.0, skyrad=SKYRAD, badpix=BADPIX, exact=True)
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(steps)]
    state = {}

    def step(k, ev=None):
        if ev:
            ev[0].record()
        mine_a, h_a = sp.aper_async(x_t, y_t, plan=plan, **kw)
        if ev:
            ev[1].record()
        # starting scale from the aperture flux (flux of a unit-scale PSF inside r=5 ~ 0.93 x its volume), sky from mmm
        vol = float(meta["gauss"][0]) * 2.0 * np.pi * float(meta["gauss"][3]) * float(meta["gauss"][4])
        sc0 = torch.clamp(torch.nan_to_num(mine_a["flux"][:, 0], nan=vol) / vol, min=1e-3)
        mine_f, h_f = sp.fit_async("pkfit_noise", None, x_t, y_t, None, FIT_RADIUS, plan=plan,
                                   sharded_inputs={"scale": sc0, "sky": mine_a["sky"]})
        if ev:
            ev[2].record()
        state["aper"], state["fit"] = h_a.result(), h_f.result()
        state["sc0_mine"] = sc0
        if ev:
            ev[3].record()

    for k in range(max(warmup, 2)):
        step(k)
    ms_total = timed(lambda k: step(k, evs[k]), steps, world, dist, torch, dev)
    ms_step = ms_total / steps
    stage = torch.tensor([np.mean([e[0].elapsed_time(e[1]) for e in evs]), np.mean([e[1].elapsed_time(e[2]) for e in evs]),
                          np.mean([e[2].elapsed_time(e[3]) for e in evs])], dtype=torch.float64, device=dev)
    allst = torch.empty((world, 3), dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_gather_into_tensor(allst, stage)
    else:
        allst[0] = stage
    out = None
    if rank == 0:
        A, F = state["aper"], state["fit"]
        nit = F["niter"]
        out = {"workload": "%dx%d fp32 mosaic+noise+u8 mask, %d stars total (one per %.1f px^2), aper exact apr=5 skyrad=[40,50] -> "
                           "recentring pkfit_noise r=5 (sky, start scale from aper), star-sharded x%d" % (size, size, ntot, size * size / ntot, world),
               "value": ntot / (ms_step * 1e-3), "unit": "stars/s", "scaling": "strong", "n_gpus": world, "steps": steps,
               "ms_per_step": ms_step, "broadcast_ms": t_bcast, "broadcast_bytes": int(size * size * 9) if world > 1 else 0,
               "field_build_s": t_gen,
               "per_rank": {"aper_ms": [round(v, 2) for v in allst[:, 0].tolist()], "pkfit_ms": [round(v, 2) for v in allst[:, 1].tolist()],
                            "exposed_gather_ms": [round(v, 2) for v in allst[:, 2].tolist()]},
               "nccl": "all_gather_into_tensor of one byte arena per stage: aper %d B/star on a side stream under the fits, "
                       "pkfit 60 B/star after them" % (8 * 6 + 1 + 8),
               "mean_niter": float(nit.double().mean().item()), "frac_nan_fit": float(torch.isnan(F["scale"]).double().mean().item()),
               "frac_bad_aper": float((A["badflag"][:, 0] != 0).double().mean().item())}
        if not a.no_parity:
            from oracle import oracle
            oracle.build()
            ns = a.cpu_sample or 20000
            idx = np.random.default_rng(11).choice(ntot, size=min(ns, ntot), replace=False)
            idx_t = torch.from_numpy(idx).to(dev)
            hf = {"image": f["image"].cpu().numpy(), "noise": f["noise"].cpu().numpy(), "mask": f["mask"].cpu().numpy(),
                  "gauss": meta["gauss"], "psf": meta["psf"], "phpadu": meta["phpadu"], "ronois": meta["ronois"], "sky": meta["sky"]}
            xs, ys = x_t[idx_t].cpu().numpy(), y_t[idx_t].cpu().numpy()
            ga, gf = to_np(A, idx_t), to_np(F, idx_t)
            # the fit's inputs are the GPU's own aper outputs (sky, start scale): the oracle is given the same numbers
            vol = float(meta["gauss"][0]) * 2.0 * np.pi * float(meta["gauss"][3]) * float(meta["gauss"][4])
            sc0 = np.maximum(np.nan_to_num(ga["flux"][:, 0], nan=vol) / vol, 1e-3)
            _, _, want = oracle_pass(hf, "config5", xs, ys, xs, ys, host_threads(), variant=oracle.PKFIT_NOISE, apr=apr5,
                                     fit_sky=ga["sky"], fit_scale=sc0)
            out["parity"] = parity_report(ga, gf, want, xs, ys, xs, ys, (size, size), ("errmag", "chi", "sharp", "niter", "scale", "x", "y"))
            del hf
    if world > 1 and not a.no_parity:
        # bitwise shard check: rank 0 recomputes a sample on its own GPU alone
        ok = None
        if rank == 0:
            from pythonphot_b200.aper import aper_device
            idx_t = torch.from_numpy(np.random.default_rng(12).choice(ntot, size=min(100000, ntot), replace=False)).to(dev)
            xa, ya = x_t[idx_t].contiguous(), y_t[idx_t].contiguous()
            a1 = aper_device(sp.image_t, xa, ya, **kw)
            vol = float(meta["gauss"][0]) * 2.0 * np.pi * float(meta["gauss"][3]) * float(meta["gauss"][4])
            sc1 = torch.clamp(torch.nan_to_num(a1["flux"][:, 0], nan=vol) / vol, min=1e-3)
            f1 = sp.engine.fit_device(_lib.PKFIT_NOISE, sc1, xa, ya, a1["sky"], FIT_RADIUS)
            ok = all(bitwise_equal(a1[k], state["aper"][k][idx_t]) for k in a1) and all(bitwise_equal(f1[k], state["fit"][k][idx_t]) for k in f1)
            out["shard_check"] = "bitwise-ok (%d stars recomputed on one GPU)" % idx_t.numel() if ok else "MISMATCH"
    del sp, state
    torch.cuda.empty_cache()
    return out


# ----------------------------------------------------------------------------- ours
def main():
    a = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if a.impl == "reference":
        return run_reference(a, rank, world)

    import torch
    import torch.distributed as dist
    import pythonphot_b200 as pp
    from pythonphot_b200 import _lib, shard
    from pythonphot_b200.aper import aper_device

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    failed = []

    if a.workload == "config5":
        c5 = run_config5(a, rank, world, dev, a.steps, a.warmup)
        if rank == 0:
            line = {"metric": "stars/sec (aper+mmm sky -> recentring pkfit_noise), BASELINE.json configs[4]", "value": c5["value"],
                    "unit": "stars/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 2), "ms_per_step": c5["ms_per_step"],
                    "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                    "impl": "ours", "config": {"workload": c5["workload"]}, "config5": c5, "gpu_launches": int(_lib.launch_count())}
            print(json.dumps(line), flush=True)
            if c5.get("parity", {}).get("mismatches") or c5.get("shard_check") == "MISMATCH":
                raise SystemExit("config5: parity / shard check failed")
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- the exposure: built on rank 0, broadcast once over NVLink (north_star), replicated in HBM
    n = a.stars
    t_b = 0.0
    if rank == 0:
        field = make_field(a)
        host = {k: torch.from_numpy(np.ascontiguousarray(field[k])).pin_memory() for k in ("image", "noise", "mask")}
    else:
        field, host = None, None
    shape = (a.size, a.size)
    tdt = torch.float64 if a.image_dtype == "float64" else torch.float32
    img_t = host["image"].to(dev, non_blocking=True) if rank == 0 else torch.empty(shape, dtype=tdt, device=dev)
    noise_t = host["noise"].to(dev, non_blocking=True) if rank == 0 else torch.empty(shape, dtype=tdt, device=dev)
    mask_t = host["mask"].to(dev, non_blocking=True) if rank == 0 else torch.empty(shape, dtype=torch.uint8, device=dev)
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        shard.broadcast_exposure([img_t, noise_t, mask_t], src=0)
        e1.record()
        torch.cuda.synchronize()
        t_b = e0.elapsed_time(e1)
        meta = shard.broadcast_header({k: field[k] for k in ("x", "y", "xfit", "yfit", "gauss", "psf", "phpadu", "ronois", "sky")}
                                      if rank == 0 else None)
        if rank != 0:
            field = meta
    # star list: `world` jittered copies of the field's list (n stars per GPU: weak scaling), known on rank 0, broadcast, then
    # sharded by image tile (pythonphot_b200/shard.py): each rank fits one chunk-interleaved Morton-ordered shard
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
    pin = lambda t: t.cpu().pin_memory()  # noqa: E731
    h_x, h_y, h_xf, h_yf = pin(d_x), pin(d_y), pin(d_xf), pin(d_yf)
    h_one, h_sky = torch.ones(n, dtype=torch.float64).pin_memory(), torch.full((n,), float(field["sky"]), dtype=torch.float64).pin_memory()
    d_one, d_sky = h_one.to(dev), h_sky.to(dev)
    pk = pp.pkfit_class(img_t, field["gauss"], field["psf"], field["ronois"], field["phpadu"], noise_t, mask_t)
    do_aper, do_fit = a.workload in ("both", "aper"), a.workload in ("both", "pkfit")

    # the kernels write their result columns straight into one byte arena per stage; at N > 1 that arena is what the
    # NCCL all-gather moves (no packing): the aper arena on a side stream, hidden by the fit kernel
    aper_arena = shard.ResultArena(shard.APER_SPEC(torch, APR.size), n, dev) if do_aper else None
    fit_arena = shard.ResultArena(shard.FIT_SPEC(torch), n, dev) if do_fit else None
    side = torch.cuda.Stream(device=dev)
    bufs = {"aper": aper_arena.views if do_aper else None, "fit": fit_arena.views if do_fit else None}
    gathered = {}

    def step(ev=None):
        h_a = h_f = None
        if do_aper:
            if ev:
                ev[0].record()
            aper_device(img_t, d_x, d_y, phpadu=field["phpadu"], apr=APR, zeropoint=25.0, skyrad=SKYRAD,
                        badpix=BADPIX, exact=True, out=aper_arena.views)
            if ev:
                ev[1].record()
            if world > 1:
                h_a = shard.gather_arena(plan, aper_arena, side_stream=side if do_fit else None)
        if do_fit:
            if ev:
                ev[2].record()
            pk.fit_device(_lib.PKFIT_NORECENT, d_one, d_xf, d_yf, d_sky, FIT_RADIUS, out=fit_arena.views)
            if ev:
                ev[3].record()
        if world > 1:  # what is left of the gather after the fits: the fit records, and the join with the side stream
            if ev:
                ev[4].record()
            if do_fit:
                h_f = shard.gather_arena(plan, fit_arena, ncols=5)  # errmag, chi, sharp, scale, niter: what the reference returns
            gathered["aper"] = h_a.result() if h_a else None
            gathered["fit"] = h_f.result() if h_f else None
            if ev:
                ev[5].record()

    for _ in range(max(a.warmup, 3)):
        step()
    torch.cuda.synchronize()
    launches0 = _lib.launch_count()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(a.steps)]
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_total = timed(lambda k: step(evs[k]), a.steps, world, dist, torch, dev)
    clocks = sampler.stop()
    launches = _lib.launch_count() - launches0
    ms_aper = float(np.mean([e[0].elapsed_time(e[1]) for e in evs])) if do_aper else 0.0
    ms_fit = float(np.mean([e[2].elapsed_time(e[3]) for e in evs])) if do_fit else 0.0
    ms_gather = float(np.mean([e[4].elapsed_time(e[5]) for e in evs])) if world > 1 else 0.0
    per_rank = None
    if world > 1:
        mine = torch.tensor([ms_aper, ms_fit, ms_gather], dtype=torch.float64, device=dev)
        allr = torch.empty((world, 3), dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(allr, mine)
        per_rank = {"aper_ms": [round(v, 2) for v in allr[:, 0].tolist()], "pkfit_ms": [round(v, 2) for v in allr[:, 1].tolist()],
                    "exposed_gather_ms": [round(v, 2) for v in allr[:, 2].tolist()]}
        ms_aper, ms_fit, ms_gather = (float(v) for v in allr.max(dim=0).values.tolist())
    ms_step = ms_total / a.steps
    value = world * n / (ms_step * 1e-3)

    # ---- one extra, untimed-for-`value` pass with the library's own CUDA events around every kernel launch (on the launch
    # stream): per-kernel average launch durations for the roofline
    _lib.profile(True)
    step()
    torch.cuda.synchronize()
    ktimes = _lib.kernel_times()
    _lib.profile(False)

    # ---- bitwise shard check (N > 1): rank 0 recomputes a sample of ALL ranks' stars on its own GPU alone
    shard_check = None
    if world > 1 and rank == 0 and not a.no_parity:
        idx_t = torch.from_numpy(np.random.default_rng(9).choice(world * n, size=min(100000, world * n), replace=False)).to(dev)
        ok = True
        if do_aper:
            a1 = aper_device(img_t, cols[0][idx_t].contiguous(), cols[1][idx_t].contiguous(), phpadu=field["phpadu"], apr=APR,
                             zeropoint=25.0, skyrad=SKYRAD, badpix=BADPIX, exact=True)
            ok &= all(bitwise_equal(a1[k], gathered["aper"][k][idx_t]) for k in a1)
        if do_fit:
            m_ = idx_t.numel()
            f1 = pk.fit_device(_lib.PKFIT_NORECENT, d_one[:m_].contiguous(), cols[2][idx_t].contiguous(), cols[3][idx_t].contiguous(),
                               d_sky[:m_].contiguous(), FIT_RADIUS)
            ok &= all(bitwise_equal(f1[k], gathered["fit"][k][idx_t]) for k in gathered["fit"])
        shard_check = ("bitwise-ok (%d stars of the gathered records recomputed on one GPU)" % idx_t.numel()) if ok else "MISMATCH"
        if not ok:
            failed.append("shard_check")

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
            if world > 1:
                # several GPUs: the exposure crosses PCIe once, on rank 0, and reaches the other GPUs over NVLink
                # (shard.broadcast_exposure); every rank then calls the public API on its own shard of
                # the star list with HOST star columns and gets HOST results back
                # band by band: rank 0's upload of band k, its NCCL broadcast, and the kernels that only need rows up to band k
                # overlap on every rank (shard.distribute_exposure)
                hdr = {"shape": shape, "image_dtype": str(tdt), "noise_dtype": str(tdt) if do_fit else None, "has_mask": bool(do_fit)}
                im_a, nz_a, mk_a = shard.distribute_exposure([np_img, np_noise if do_fit else None, np_mask if do_fit else None],
                                                             header=hdr, src=0, device=dev)
                if rank == 0:
                    h2d += np_img.nbytes + (np_noise.nbytes + np_mask.nbytes if do_fit else 0)
            else:
                im_a, nz_a, mk_a = np_img, np_noise, np_mask
                h2d += np_img.nbytes + (np_noise.nbytes + np_mask.nbytes if do_fit else 0)
            pkh = None
            if do_fit:
                # the reference's own order of construction is free: pkfit_class first, so that its uploads (image, then
                # noise and mask, on a copy stream) run while aper computes; aper is handed the SAME device image
                pkh = pp.pkfit_class(im_a, field["gauss"], field["psf"], field["ronois"], field["phpadu"], nz_a, mk_a)
                h2d += field["psf"].nbytes
            if do_aper:
                r = pp.aper.aper(pkh.image_device() if pkh is not None else im_a, nx, ny, phpadu=field["phpadu"], apr=APR,
                                 zeropoint=25.0, skyrad=list(SKYRAD), badpix=list(BADPIX), exact=True, verbose=False)
                h2d += nx.nbytes + ny.nbytes
                d2h += sum(np.asarray(v).nbytes for v in r[:6]) + r[6].size + 8 * n  # uint8 flags + nsky/mmm_iter
            if do_fit:
                r = pkh.pkfit_norecent_noise(none, nxf, nyf, nsky, FIT_RADIUS)
                h2d += 4 * nxf.nbytes
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
               "path": ("pk = pkfit_class(numpy image, noise, mask) [one upload of each, on a copy stream]; "
                        "aper.aper(pk.image_device(), numpy x, y) -> numpy results; pk.pkfit_norecent_noise(numpy lists) -> numpy results"
                        if world == 1 else
                        "rank 0 uploads the exposure in row bands and NCCL-broadcasts each band as it lands (shard.distribute_exposure); "
                        "every rank calls pkfit_class + aper.aper + pkfit_norecent_noise on its shard with numpy star columns, numpy "
                        "results, its kernels starting on the bands already there; byte counts are rank 0's")}

    # ---- config 5 (N > 1): the multi-GPU configuration north_star names, strong scaling
    c5 = None
    if world > 1 and not a.no_config5 and a.workload == "both":
        torch.cuda.empty_cache()
        c5 = run_config5(a, rank, world, dev, steps=max(2, min(a.steps, 3)), warmup=2)

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
                              "achieved_gbs": n * b_aper / (ms_aper * 1e-3) / 1e9,
                              "mean_mmm_iter": float(bufs["aper"]["mmm_iter"].double().mean().item())}
        if do_fit:
            nit = bufs["fit"]["niter"]
            stages["pkfit"] = {"ms": ms_fit, "stars_per_s": n / (ms_fit * 1e-3), "job_stars_per_s": world * n / (ms_fit * 1e-3), "alg_bytes_per_star": b_fit,
                               "achieved_gbs": n * b_fit / (ms_fit * 1e-3) / 1e9,
                               "mean_niter": float(nit.double().mean().item()),
                               "frac_niter_25": float((nit == 25).double().mean().item())}
        if world > 1:
            stages["per_rank"] = per_rank
            nb_a = aper_arena.nbytes if do_aper else 0
            nb_f = 36 * n if do_fit else 0  # errmag, chi, sharp, scale (fp64) + niter (int32): the gathered prefix of the fit arena
            stages["gather"] = {"ms": ms_gather, "what": "EXPOSED part: all-gather + un-permute of the fit records, join with the side stream",
                                "nccl": "all_gather_into_tensor x2 per step: aper arena %d B/rank on a side stream under the fit kernel, "
                                        "fit arena %d B/rank after it" % (nb_a, nb_f)}
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
                "data": "synthetic", "impl": "ours", "config": config_dict(a, world), "broadcast_ms": t_b,
                "build_flags": _lib.build_flags(),
                "stages": stages, "kernels": kernels, "roofline": roofline, "clocks": clocks, "gpu_launches": int(launches), "e2e": e2e}
        if shard_check:
            line["shard_check"] = shard_check
        if c5 is not None:
            line["config5"] = c5
            if c5.get("parity", {}).get("mismatches") or c5.get("shard_check") == "MISMATCH":
                failed.append("config5")
        # ---- CPU legs: the oracle on a sample of the SAME exposure and star list -- timed (cpu_baseline, N = 1) and compared
        # with what the timed GPU steps left in their result buffers (parity, every N)
        want_base = world == 1 and not a.no_cpu_baseline
        if want_base or not a.no_parity:
            from oracle import oracle
            oracle.build()
            if world == 1:
                ns = a.cpu_sample or 400000
                info, idx, want = cpu_baseline(field, a, ns, keep=True)
                if want_base:
                    line["cpu_baseline"] = info
                xs, ys, xfs, yfs = field["x"][idx], field["y"][idx], field["xfit"][idx], field["yfit"][idx]
                src_a, src_f = bufs["aper"], bufs["fit"]
            else:  # the gathered records of every rank's stars, in caller order
                ns = a.cpu_sample or 50000
                idx = np.random.default_rng(7).choice(world * n, size=min(ns, world * n), replace=False)
                xs, ys, xfs, yfs = (c[idx] for c in full)
                _, _, want = oracle_pass(field, a.workload, xs, ys, xfs, yfs, host_threads())
                src_a, src_f = gathered.get("aper"), gathered.get("fit")
            if not a.no_parity:
                idx_t = torch.from_numpy(idx).to(dev)
                if world == 1:  # results land in the order of this rank's shard (= plan order), not caller order
                    idx_t = plan._index_t("inverse", 0, dev)[idx_t]
                rep = parity_report(to_np(src_a, idx_t) if do_aper else None, to_np(src_f, idx_t) if do_fit else None, want,
                                    xs, ys, xfs, yfs, shape,
                                    ("errmag", "chi", "sharp", "niter", "scale") + (("chi2",) if world == 1 else ()))
                line["parity"] = rep
                if rep["mismatches"]:
                    failed.append("parity")
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if failed:
        raise SystemExit("bench.py: %s failed -- the timed results do not match the oracle / the single-GPU run" % ", ".join(failed))


if __name__ == "__main__":
    main()
