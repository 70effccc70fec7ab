"""Star sharding: host logic on CPU, world_size 2 over gloo; the CUDA path at world 1 / 2 on the box."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from pythonphot_b200 import shard  # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_compute(cols):
    """Stands in for the CUDA kernels: a per-star record that depends on that star only."""
    x, y = cols["x"], cols["y"]
    flux = torch.stack([x * 3.0 + y, x - y * 0.5, x * y], dim=1)
    return {"flux": flux, "sky": x + 0.25 * y, "niter": (x.abs() % 25).to(torch.int32),
            "badflag": ((x + y) > 500).to(torch.uint8).unsqueeze(1).repeat(1, 3)}


def _worker(rank, world, port, n, out_dir):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(5)
        x, y = rng.uniform(-5, 1030, n), rng.uniform(-5, 1030, n)
        if n > 3:
            x[1], y[2] = np.nan, np.inf
        # only rank 0 "has" the star list and the exposure
        cols = shard.broadcast_columns([x, y] if rank == 0 else [None, None], n, "cpu", src=0)
        img = torch.arange(12, dtype=torch.float32).reshape(3, 4) if rank == 0 else torch.empty(3, 4)
        mask = torch.ones(3, 4, dtype=torch.uint8) if rank == 0 else torch.empty(3, 4, dtype=torch.uint8)
        shard.broadcast_exposure([img, None, mask], src=0)
        assert torch.equal(img, torch.arange(12, dtype=torch.float32).reshape(3, 4)) and bool(mask.all())
        plan = shard.ShardPlan.for_stars(cols[0].numpy(), cols[1].numpy(), 1024, 1024, world)
        assert len(plan.indices(rank)) == plan.m
        res = shard.run_sharded(plan, {"x": cols[0], "y": cols[1]}, _fake_compute)
        torch.save({k: v for k, v in res.items()}, os.path.join(out_dir, "r%d.pt" % rank))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [1, 7, 1001])
def test_run_sharded_world2_equals_single_process(tmp_path, n):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n, str(tmp_path)), nprocs=2, join=True)
    rng = np.random.default_rng(5)
    x, y = rng.uniform(-5, 1030, n), rng.uniform(-5, 1030, n)
    if n > 3:
        x[1], y[2] = np.nan, np.inf
    want = _fake_compute({"x": torch.from_numpy(x), "y": torch.from_numpy(y)})
    for r in range(2):
        got = torch.load(os.path.join(str(tmp_path), "r%d.pt" % r))
        assert set(got) == set(want)
        for k in want:
            assert got[k].dtype == want[k].dtype and got[k].shape == want[k].shape, k
            # bit for bit, NaN payloads included
            assert torch.equal(got[k].contiguous().view(torch.uint8), want[k].contiguous().view(torch.uint8)), (k, r)


def test_plan_covers_every_star_once_and_balances():
    rng = np.random.default_rng(1)
    for n, world in [(10, 1), (10, 3), (8, 8), (5, 8), (100001, 8)]:
        x, y = rng.uniform(0, 4096, n), rng.uniform(0, 4096, n)
        plan = shard.ShardPlan.for_stars(x, y, 4096, 4096, world)
        allidx = np.concatenate([plan.indices(r) for r in range(world)])
        assert allidx.size == plan.m * world and plan.m == -(-n // world)
        assert np.array_equal(np.sort(np.unique(allidx)), np.arange(n))
        assert np.array_equal(allidx[plan.inverse], np.arange(n))


def test_morton_order_groups_tiles():
    # stars of one 64-px tile are contiguous in the visiting order
    rng = np.random.default_rng(2)
    x, y = rng.uniform(0, 512, 5000), rng.uniform(0, 512, 5000)
    p = shard.morton_order(x, y, 512, 512)
    tiles = (y[p].astype(int) >> 6) * 8 + (x[p].astype(int) >> 6)
    change = np.flatnonzero(np.diff(tiles) != 0).size
    assert change == np.unique(tiles).size - 1
    # stable: equal tiles keep the caller's relative order
    for t in np.unique(tiles)[:5]:
        idx = p[tiles == t]
        assert np.all(np.diff(idx) > 0)


def test_pack_unpack_roundtrip():
    cols = {"a": torch.randn(9, 3, dtype=torch.float64), "b": torch.arange(9, dtype=torch.int32),
            "c": torch.randint(0, 2, (9, 3), dtype=torch.uint8)}
    cols["d"] = torch.randn(9, dtype=torch.float64)
    packed, layout = shard.pack_records(cols)
    assert {dt: tuple(b.shape) for dt, b in packed.items()} == {torch.float64: (9, 4), torch.int32: (9, 1), torch.uint8: (9, 3)}
    back = shard.unpack_records(packed, layout)
    for k in cols:
        assert torch.equal(back[k], cols[k])


# ----------------------------------------------------------------------------- CUDA path
def _gpu_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    try:
        from pythonphot_b200 import synth
        f = synth.make_field(768, 768, 3000, seed=11, dtype=np.float32)
        n = f["x"].size
        sp = shard.ShardedPhotometry(f["image"] if rank == 0 else None, f["gauss"], f["psf"], f["ronois"], f["phpadu"],
                                     f["noise"] if rank == 0 else None, f["mask"] if rank == 0 else None, shape=(768, 768))
        a = sp.aper(f["x"], f["y"], phpadu=f["phpadu"], apr=np.array([3., 5., 8.]), skyrad=[40, 50],
                    badpix=[-12000, 60000], exact=True)
        p = sp.fit("pkfit_noise", np.ones(n), f["xfit"], f["yfit"], a["sky"], 5.0)
        torch.save({"aper": {k: v.cpu() for k, v in a.items()}, "fit": {k: v.cpu() for k, v in p.items()}},
                   os.path.join(out_dir, "g%d.pt" % rank))
    finally:
        dist.destroy_process_group()


def _single_gpu_reference():
    import pythonphot_b200 as pp
    from pythonphot_b200 import synth, _lib
    from pythonphot_b200.aper import aper_device
    f = synth.make_field(768, 768, 3000, seed=11, dtype=np.float32)
    n = f["x"].size
    dev = torch.device("cuda", 0)
    pk = pp.pkfit_class(f["image"], f["gauss"], f["psf"], f["ronois"], f["phpadu"], f["noise"], f["mask"], device=dev)
    a = aper_device(pk._img_t, _lib.to_device_f64(f["x"], dev), _lib.to_device_f64(f["y"], dev), phpadu=f["phpadu"],
                    apr=np.array([3., 5., 8.]), skyrad=[40, 50], badpix=[-12000, 60000], exact=True)
    p = pk.fit_device(_lib.PKFIT_NOISE, _lib.to_device_f64(np.ones(n), dev), _lib.to_device_f64(f["xfit"], dev),
                      _lib.to_device_f64(f["yfit"], dev), a["sky"], 5.0)
    return {"aper": {k: v.cpu() for k, v in a.items()}, "fit": {k: v.cpu() for k, v in p.items()}}


def _assert_bitwise(got, want):
    for stage in want:
        for k in want[stage]:
            g, w = got[stage][k].contiguous(), want[stage][k].contiguous()
            assert g.dtype == w.dtype and g.shape == w.shape, (stage, k)
            assert torch.equal(g.view(torch.uint8), w.view(torch.uint8)), "%s.%s differs from the single-GPU run" % (stage, k)


@pytest.mark.gpu
def test_sharded_world1_equals_direct_call(tmp_path):
    """No process group: ShardedPhotometry degenerates to the tile-ordered single-GPU pass."""
    _gpu_worker_single(str(tmp_path))
    _assert_bitwise(torch.load(os.path.join(str(tmp_path), "g0.pt")), _single_gpu_reference())


def _gpu_worker_single(out_dir):
    from pythonphot_b200 import synth
    f = synth.make_field(768, 768, 3000, seed=11, dtype=np.float32)
    n = f["x"].size
    sp = shard.ShardedPhotometry(f["image"], f["gauss"], f["psf"], f["ronois"], f["phpadu"], f["noise"], f["mask"])
    a = sp.aper(f["x"], f["y"], phpadu=f["phpadu"], apr=np.array([3., 5., 8.]), skyrad=[40, 50], badpix=[-12000, 60000], exact=True)
    p = sp.fit("pkfit_noise", np.ones(n), f["xfit"], f["yfit"], a["sky"], 5.0)
    torch.save({"aper": {k: v.cpu() for k, v in a.items()}, "fit": {k: v.cpu() for k, v in p.items()}}, os.path.join(out_dir, "g0.pt"))


@pytest.mark.gpu
def test_sharded_two_gpus_bit_identical_to_one(tmp_path):
    """SURVEY.md 8e correctness check: N-GPU output equals 1-GPU output bit for bit after un-permuting."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_gpu_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    want = _single_gpu_reference()
    for r in range(2):
        _assert_bitwise(torch.load(os.path.join(str(tmp_path), "g%d.pt" % r)), want)


def _gpu_worker_banded(rank, world, port, out_dir):
    """Exposure distributed band by band (upload on rank 0 + NCCL broadcast per band), consumed through DeviceImage."""
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    try:
        import pythonphot_b200 as pp
        from pythonphot_b200 import synth
        f = synth.make_field(768, 768, 3000, seed=11, dtype=np.float32)
        n = f["x"].size
        arrays = [f["image"], f["noise"], f["mask"]] if rank == 0 else [None, None, None]
        dimg, dnz, dmk = shard.distribute_exposure(arrays, src=0, nbands=7)
        assert len(dimg.events) == 7 and dmk.tensor.dtype == torch.uint8
        pk = pp.pkfit_class(dimg, f["gauss"], f["psf"], f["ronois"], f["phpadu"], dnz, dmk)
        a = pp.aper.aper(pk.image_device(), f["x"], f["y"], phpadu=f["phpadu"], apr=np.array([3., 5., 8.]), skyrad=[40, 50],
                         badpix=[-12000, 60000], exact=True, verbose=False)
        p = pk.pkfit_noise(np.ones(n), f["xfit"], f["yfit"], a[4], 5.0, xyout=True)
        torch.save({"aper": [np.asarray(v) for v in a[:7]], "fit": [np.asarray(v) for v in p]}, os.path.join(out_dir, "b%d.pt" % rank))
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
def test_banded_exposure_distribution_two_gpus(tmp_path):
    """Every rank computes the full star list on ITS copy of the exposure, received band by band: all copies must give what
    the plain single-GPU call on host arrays gives, bit for bit."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    import torch.multiprocessing as mp
    import pythonphot_b200 as pp
    from pythonphot_b200 import synth
    port = _free_port()
    mp.spawn(_gpu_worker_banded, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    f = synth.make_field(768, 768, 3000, seed=11, dtype=np.float32)
    n = f["x"].size
    a = pp.aper.aper(f["image"], f["x"], f["y"], phpadu=f["phpadu"], apr=np.array([3., 5., 8.]), skyrad=[40, 50],
                     badpix=[-12000, 60000], exact=True, verbose=False)
    pk = pp.pkfit_class(f["image"], f["gauss"], f["psf"], f["ronois"], f["phpadu"], f["noise"], f["mask"])
    p = pk.pkfit_noise(np.ones(n), f["xfit"], f["yfit"], a[4], 5.0, xyout=True)
    for r in range(2):
        got = torch.load(os.path.join(str(tmp_path), "b%d.pt" % r), weights_only=False)
        for u, v in zip(got["aper"], a[:7]):
            assert np.array_equal(np.nan_to_num(u, nan=-1.0), np.nan_to_num(np.asarray(v), nan=-1.0))
        for u, v in zip(got["fit"], p):
            assert np.array_equal(np.nan_to_num(u, nan=-1.0), np.nan_to_num(np.asarray(v), nan=-1.0))
