"""Star-sharded photometry over several GPUs of one node.

Stars are independent units (aper.py:217 and README.md:73-75 loop over them one at a time), so the
path shards by star list and needs exactly two collectives, both outside the kernels:

  * the exposure (image, noise image, mask) is broadcast ONCE from the source rank over
    NVLink/NVSwitch (``broadcast_exposure``) and stays replicated in every GPU's HBM;
  * each rank runs the ordinary single-GPU kernels on its contiguous shard of the tile-ordered star
    list and the fixed-width per-star result records are all-gathered (``ShardPlan.gather``).

The kernels write their result columns straight into one byte arena per call (``ResultArena``: the columns are typed views
of one flat buffer), so a pass costs ONE all-gather of that buffer -- no packing step -- and one indexed copy per column into
caller order; ``gather_arena(..., side_stream=...)`` runs both on a side stream so that the next kernel (the PSF fits after
``aper``) hides them.  There is no collective inside a kernel, hence nothing to fuse.  The plumbing below is
``torch.distributed`` only (NCCL on the GPU box; the same code runs under ``gloo`` with CPU tensors,
which is how tests/test_shard_gloo.py covers it at world_size 2).  The N-GPU result equals the
1-GPU result bit for bit after un-permuting, because every star's arithmetic is independent of
which rank, CTA or batch it lands in.
"""
import numpy as np

TILE_SHIFT = 6  # 64-pixel tiles, as SURVEY.md 8e


def _dist():
    import torch.distributed as dist
    return dist


def world_info(group=None):
    """(rank, world) of the default/explicit process group, (0, 1) when torch.distributed is not up."""
    dist = _dist()
    if not (dist.is_available() and dist.is_initialized()):
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def _spread_bits(v):
    """Interleave helper: 16-bit integers -> their bits at even positions of a 32-bit word."""
    v = v.astype(np.uint32) & np.uint32(0xFFFF)
    v = (v | (v << np.uint32(8))) & np.uint32(0x00FF00FF)
    v = (v | (v << np.uint32(4))) & np.uint32(0x0F0F0F0F)
    v = (v | (v << np.uint32(2))) & np.uint32(0x33333333)
    v = (v | (v << np.uint32(1))) & np.uint32(0x55555555)
    return v


def morton_order(x, y, nrow, ncol, shift=TILE_SHIFT):
    """Stable permutation that sorts stars by the Morton code of their 64x64-pixel image tile, so that a
    contiguous shard covers a compact patch of the frame (stamp reads of one rank share L2 lines).
    NaN / off-frame positions are clamped into the frame first (they still get a slot: the kernels report
    them in band, aper.py:221-233)."""
    xv = np.nan_to_num(np.asarray(x, dtype=np.float64), nan=0.0, posinf=float(ncol - 1), neginf=0.0)
    yv = np.nan_to_num(np.asarray(y, dtype=np.float64), nan=0.0, posinf=float(nrow - 1), neginf=0.0)
    tx = np.clip(xv, 0, max(ncol - 1, 0)).astype(np.int64) >> shift
    ty = np.clip(yv, 0, max(nrow - 1, 0)).astype(np.int64) >> shift
    code = _spread_bits(tx) | (_spread_bits(ty) << np.uint32(1))
    return np.argsort(code, kind="stable")


class ShardPlan(object):
    """Which stars each rank processes, and how the gathered records go back into caller order.

    ``perm`` (length n) is the visiting order; it is cut into ``world`` contiguous shards of ``m =
    ceil(n / world)`` stars, the last shard(s) padded by repeating the final star (cheap, discarded
    after the gather) so that every rank runs the same amount of work and the all-gather is regular."""

    def __init__(self, n, world, perm=None):
        self.n, self.world = int(n), int(world)
        self.m = (self.n + self.world - 1) // self.world if self.n else 0
        perm = np.arange(self.n, dtype=np.int64) if perm is None else np.asarray(perm, dtype=np.int64)
        assert perm.shape == (self.n,)
        pad = self.m * self.world - self.n
        self.padded = np.concatenate([perm, np.full(pad, perm[-1] if self.n else 0, dtype=np.int64)]) if pad else perm
        # position of star s in the gathered (rank-major) array
        self.inverse = np.empty(self.n, dtype=np.int64)
        self.inverse[perm] = np.arange(self.n, dtype=np.int64)
        self._dev = {}  # (which, device) -> index tensor, uploaded once

    def _index_t(self, which, rank, device):
        import torch
        key = (which, rank, str(device))
        if key not in self._dev:
            if which == "inverse":
                idx = self.inverse
            elif which == "rank_of":  # which rank's shard holds star s, and where in it
                idx = self.inverse // max(self.m, 1)
            elif which == "pos_of":
                idx = self.inverse % max(self.m, 1)
            else:
                idx = self.indices(rank)
            self._dev[key] = torch.from_numpy(np.ascontiguousarray(idx)).to(device)
        return self._dev[key]

    def _flat_index(self, es, k, nbytes, device):
        """int64 [n * k]: where component j of star s sits in the gathered [world, nbytes] buffer viewed as elements of
        ``es`` bytes, relative to the start of its column (cached per element size, width and arena size)."""
        import torch
        key = ("flat", es, k, nbytes, str(device))
        if key not in self._dev:
            base = (self.inverse // max(self.m, 1)) * (nbytes // es) + (self.inverse % max(self.m, 1)) * k
            idx = base if k == 1 else (base[:, None] + np.arange(k, dtype=np.int64)[None, :]).reshape(-1)
            self._dev[key] = torch.from_numpy(np.ascontiguousarray(idx)).to(device)
        return self._dev[key]

    @classmethod
    def for_stars(cls, x, y, nrow, ncol, world, chunk=4096):
        """Tile-ordered stars dealt to the ranks in chunks of ``chunk`` consecutive stars (chunk c -> rank c mod world):
        a chunk keeps the L2 locality of its image tiles, and every rank sees the same mix of crowded, sparse and
        frame-edge regions, so per-rank run times stay level (one contiguous Morton patch per rank measured 20 % apart
        on 8 GPUs)."""
        perm = morton_order(x, y, nrow, ncol)
        if world > 1 and chunk > 0:
            owner = (np.arange(perm.size) // chunk) % world
            perm = perm[np.argsort(owner, kind="stable")]
        return cls(len(x), world, perm)

    def indices(self, rank):
        """Star indices (into the caller's list) of ``rank``'s shard, padding included."""
        return self.padded[rank * self.m:(rank + 1) * self.m]

    def take(self, column, rank):
        """``rank``'s shard of a star-list column (numpy array or torch tensor, first axis = stars)."""
        idx = self.indices(rank)
        if isinstance(column, np.ndarray):
            return np.ascontiguousarray(column[idx])
        return column.index_select(0, self._index_t("shard", rank, column.device)).contiguous()

    def gather(self, record, group=None):
        """All-gather one per-star record tensor ``[m, ...]`` (same shape and dtype on every rank) and
        return it as ``[n, ...]`` in the caller's star order, on every rank."""
        import torch
        dist = _dist()
        rank, world = world_info(group)
        assert world == self.world, "plan was made for %d ranks, group has %d" % (self.world, world)
        assert record.shape[0] == self.m, "record has %d rows, shard has %d" % (record.shape[0], self.m)
        record = record.contiguous()
        if world == 1:
            full = record
        else:
            full = torch.empty((world * self.m,) + tuple(record.shape[1:]), dtype=record.dtype, device=record.device)
            dist.all_gather_into_tensor(full, record, group=group)
        return full.index_select(0, self._index_t("inverse", 0, full.device))


def gather_arena(plan, arena, group=None, side_stream=None, ncols=None):
    """All-gather a ``ResultArena`` (one NCCL ``all_gather_into_tensor`` of ``arena.nbytes`` per rank) and copy every column
    into caller order.  With ``side_stream`` (CUDA) both run on that stream, after the work already queued on the current
    one; the returned ``GatherHandle.result()`` joins.  ``ncols`` gathers only the first columns of the arena's spec (a
    prefix of the flat buffer).  Works on CPU tensors under gloo (tests)."""
    return _gather_arena(plan, arena, group=group, side_stream=side_stream, ncols=ncols)


def pack_records(columns):
    """Dict of per-star tensors ``[m]`` / ``[m, k]`` -> one ``[m, width]`` tensor PER DTYPE plus the layout needed
    to unpack, so that a pass costs one all-gather and one un-permute per dtype (fp64, int32, uint8) instead of
    one per column, and every copy moves whole typed words."""
    import torch
    groups, layout = {}, []
    for name, t in columns.items():
        t2 = t.reshape(t.shape[0], -1)
        parts = groups.setdefault(t.dtype, [])
        start = sum(p.shape[1] for p in parts)
        parts.append(t2)
        layout.append((name, t.dtype, tuple(t.shape[1:]), start, t2.shape[1]))
    packed = {dt: (parts[0].contiguous() if len(parts) == 1 else torch.cat(parts, dim=1)) for dt, parts in groups.items()}
    return packed, layout


def unpack_records(packed, layout):
    out = {}
    for name, dtype, tail, start, width in layout:
        block = packed[dtype]
        out[name] = block[:, start:start + width].reshape((block.shape[0],) + tail)
    return out


_NP_DTYPES = {"torch.float64": np.float64, "torch.float32": np.float32, "torch.int32": np.int32, "torch.uint8": np.uint8,
              "torch.int64": np.int64}


class ResultArena(object):
    """Per-star result columns of one shard as typed views of ONE flat byte buffer.

    ``spec`` = [(name, torch dtype, tail shape)], ``m`` = stars in the shard.  The kernels are handed ``views`` as their
    output tensors (``aper_device(out=...)``, ``PkfitEngine.fit_device(out=...)``), so nothing has to be packed before
    the all-gather: ``flat`` is what travels.  Column offsets are 256-byte aligned."""

    def __init__(self, spec, m, device):
        import torch
        self.spec = [(n_, dt, tuple(tail)) for n_, dt, tail in spec]
        self.m, self.device = int(m), device
        self.offsets, off = {}, 0
        for name, dt, tail in self.spec:
            nb = self.m * int(np.prod(tail, dtype=np.int64)) * torch.empty((), dtype=dt).element_size()
            self.offsets[name] = (off, nb)
            off += (nb + 255) // 256 * 256
        self.nbytes = max(off, 256)
        self.flat = torch.zeros(self.nbytes, dtype=torch.uint8, device=device)
        self.views = {}
        for name, dt, tail in self.spec:
            o, nb = self.offsets[name]
            self.views[name] = self.flat[o:o + nb].view(dt).view((self.m,) + tail)

    @classmethod
    def like(cls, results, device=None):
        """Arena with one column per tensor of ``results`` (dict name -> [m, ...] tensor)."""
        first = next(iter(results.values()))
        spec = [(k, v.dtype, tuple(v.shape[1:])) for k, v in results.items()]
        return cls(spec, first.shape[0], device if device is not None else first.device)

    def to_host(self):
        """The whole arena in ONE device->host copy through a page-locked block; returns {column: numpy view of that block}.
        (Nine separate column copies of a 1 M-star ``aper`` result took 5-9 ms on B200, one 144 MB copy 2.7 ms.)"""
        import torch
        h = torch.empty(self.nbytes, dtype=torch.uint8, pin_memory=True)
        h.copy_(self.flat, non_blocking=True)
        torch.cuda.current_stream(self.flat.device).synchronize()
        hn = h.numpy()
        out = {}
        for name, dt, tail in self.spec:
            o, nb = self.offsets[name]
            out[name] = hn[o:o + nb].view(_NP_DTYPES[str(dt)]).reshape((self.m,) + tail)
        return out

    def store(self, results):
        """Copy result tensors that are not already this arena's views into it."""
        for k, v in results.items():
            if v.data_ptr() != self.views[k].data_ptr():
                self.views[k].copy_(v)
        return self.views


class GatherHandle(object):
    """An all-gather + un-permute in flight; ``result()`` makes the current stream wait for it and returns the dict of
    full-length tensors in the caller's star order."""

    def __init__(self, outs, side_stream, device):
        self._outs, self._side, self._device = outs, side_stream, device

    def result(self):
        if self._side is not None:
            import torch
            torch.cuda.current_stream(self._device).wait_stream(self._side)
            self._side = None
        return self._outs


def _gather_arena(plan, arena, group=None, side_stream=None, ncols=None):
    import torch
    dist = _dist()
    rank, world = world_info(group)
    assert world == plan.world, "plan was made for %d ranks, group has %d" % (plan.world, world)
    assert arena.m == plan.m, "arena has %d rows, shard has %d" % (arena.m, plan.m)
    dev = arena.flat.device
    use_side = side_stream is not None and dev.type == "cuda"
    # buffers are allocated on the caller's stream; the side stream only uses them between wait_stream pairs
    spec = arena.spec if ncols is None else arena.spec[:ncols]
    last_off, last_nb = arena.offsets[spec[-1][0]]
    nbytes = arena.nbytes if ncols is None else (last_off + last_nb + 255) // 256 * 256  # the prefix that holds `spec`
    flat = arena.flat[:nbytes]
    full = flat.reshape(1, -1) if world == 1 else torch.empty((world, nbytes), dtype=torch.uint8, device=dev)
    outs = {name: torch.empty((plan.n,) + tail, dtype=dt, device=dev) for name, dt, tail in spec}

    def work():
        if world > 1:
            dist.all_gather_into_tensor(full.reshape(-1), flat, group=group)
        for name, dt, tail in spec:
            o, nb = arena.offsets[name]
            es = outs[name].element_size()
            k = int(np.prod(tail, dtype=np.int64))
            typed = full.reshape(-1).view(dt)
            if plan.n:  # element (s, j) <- flat element rank_of[s] * (nbytes / es) + pos_of[s] * k + j of the column: one gather
                torch.take(typed[o // es:], plan._flat_index(es, k, nbytes, dev), out=outs[name].reshape(-1))
    if use_side:
        cur = torch.cuda.current_stream(dev)
        side_stream.wait_stream(cur)
        with torch.cuda.stream(side_stream):
            work()
        return GatherHandle(outs, side_stream, dev)
    work()
    return GatherHandle(outs, None, dev)


def broadcast_exposure(tensors, src=0, group=None):
    """Broadcast the exposure tensors (image, noise, mask; already allocated with the right shape and dtype
    on every rank) from ``src``.  One NCCL broadcast per tensor over NVLink; returns the tensors."""
    dist = _dist()
    _, world = world_info(group)
    if world > 1:
        for t in tensors:
            if t is not None:
                dist.broadcast(t, src=src, group=group)
    return tensors


def distribute_exposure(arrays, header=None, src=0, group=None, device=None, nbands=16):
    """Host exposure on ``src`` -> one ``_lib.DeviceImage`` per array on EVERY rank, pipelined band by band: ``src`` uploads
    band k on its copy stream, the NCCL broadcast of band k follows on a communication stream as soon as that band is
    resident, and the event of band k fires on each rank when ITS copy of rows [0, (k+1)*band_rows) is complete -- so
    ``aper.aper(DeviceImage, ...)`` starts on every GPU while the rest of the exposure is still crossing PCIe and NVLink.

    ``arrays`` = [image, noise, mask] (host arrays or CUDA tensors; entries may be None) on ``src``, anything elsewhere;
    ``header`` = ``exposure_header``-style dict known on every rank (else ``src`` broadcasts it first: a host-synchronous
    object broadcast).  The arrays are sent one after the other, so a later one's events imply the earlier ones.
    Returns [image, noise, mask] as ``DeviceImage`` (None where the source has none)."""
    from . import _lib
    torch = _lib.torch_cuda()
    dist = _dist()
    rank, world = world_info(group)
    dev = _lib.device_of(device)
    ups = [None, None, None]
    if rank == src:
        ups[0] = _lib.upload_image_banded(arrays[0], dev, nbands=nbands, min_bytes=0)
        ups[1] = _lib.upload_image_banded(arrays[1], dev, nbands=nbands, min_bytes=0) if arrays[1] is not None else None
        ups[2] = _lib.upload_mask_banded(arrays[2], dev, nbands=nbands, min_bytes=0) if arrays[2] is not None else None
    if header is None:
        header = broadcast_header(exposure_header(*[u.tensor if u is not None else None for u in ups]) if rank == src else None,
                                  src=src, group=group)
    if world == 1:
        return ups
    dts = {"torch.float32": torch.float32, "torch.float64": torch.float64}
    shape = tuple(header["shape"])
    kinds = [dts[header["image_dtype"]], dts[header["noise_dtype"]] if header["noise_dtype"] else None,
             torch.uint8 if header["has_mask"] else None]
    rows = (shape[0] + nbands - 1) // nbands
    cur, comm = torch.cuda.current_stream(dev), _comm_stream(dev)
    comm.wait_stream(cur)
    out = []
    for up, dt in zip(ups, kinds):
        if dt is None:
            out.append(None)
            continue
        t = up.tensor if up is not None else torch.empty(shape, dtype=dt, device=dev)
        events = []
        with torch.cuda.stream(comm):
            for k, r0 in enumerate(range(0, shape[0], rows)):
                if up is not None and up.events:
                    comm.wait_event(up.events[min(k, len(up.events) - 1)])
                elif up is not None and k == 0:
                    comm.wait_stream(cur)  # a source tensor that was produced on the caller's stream
                dist.broadcast(t[r0:r0 + rows], src=src, group=group)
                ev = torch.cuda.Event()
                ev.record(comm)
                events.append(ev)
        t.record_stream(comm)
        out.append(_lib.DeviceImage(t, events, rows))
    return out


_comm_streams = {}


def _comm_stream(device):
    import torch
    key = str(device)
    if key not in _comm_streams:
        _comm_streams[key] = torch.cuda.Stream(device=device)
    return _comm_streams[key]


def broadcast_columns(columns, n, device, src=0, group=None):
    """Star-list columns (fp64) known on ``src`` -> a ``[ncol, n]`` fp64 tensor on every rank."""
    import torch
    dist = _dist()
    rank, world = world_info(group)
    if rank == src:
        if all(isinstance(c, torch.Tensor) for c in columns):
            buf = torch.stack([c.to(device=device, dtype=torch.float64).reshape(-1) for c in columns]).contiguous()
        else:
            buf = torch.from_numpy(np.ascontiguousarray(np.stack([np.asarray(c, dtype=np.float64).reshape(-1) for c in columns]))).to(device)
        assert buf.shape[1] == n
    else:
        buf = torch.empty((len(columns), n), dtype=torch.float64, device=device)
    if world > 1:
        dist.broadcast(buf, src=src, group=group)
    return buf


def run_sharded(plan, columns, compute, group=None):
    """The generic sharded pass: ``columns`` = dict of full-length star-list columns (torch tensors on this
    rank's device, identical on every rank); ``compute(shard_columns) -> dict of per-star result tensors``
    runs this rank's shard (on the GPU box: the CUDA kernels); the records go through a ``ResultArena`` (one all-gather)
    and come back as a dict of full-length tensors in the caller's star order."""
    rank, _ = world_info(group)
    shard = {k: plan.take(v, rank) for k, v in columns.items()}
    res = compute(shard)
    arena = ResultArena.like(res)
    arena.store(res)
    return gather_arena(plan, arena, group=group).result()


APER_SPEC = lambda torch, naper: [("mag", torch.float64, (naper,)), ("magerr", torch.float64, (naper,)),  # noqa: E731
                                  ("flux", torch.float64, (naper,)), ("fluxerr", torch.float64, (naper,)),
                                  ("sky", torch.float64, ()), ("skyerr", torch.float64, ()),
                                  ("badflag", torch.uint8, (naper,)), ("nsky", torch.int32, ()), ("mmm_iter", torch.int32, ())]
# the reference's own five return values first: a caller that does not need x, y, chi2 gathers only that prefix of the arena
FIT_SPEC = lambda torch: [(k, torch.float64, ()) for k in ("errmag", "chi", "sharp", "scale")] + [("niter", torch.int32, ())] + \
    [(k, torch.float64, ()) for k in ("x", "y", "chi2")]  # noqa: E731


def exposure_header(image_t, noise_t, mask_t):
    """What the other ranks need to allocate their replica of the exposure."""
    return {"shape": tuple(image_t.shape), "image_dtype": str(image_t.dtype), "noise_dtype": str(noise_t.dtype) if noise_t is not None else None,
            "has_mask": mask_t is not None}


def broadcast_header(header, src=0, group=None):
    """``header`` (a small picklable object known on ``src``) -> the same object on every rank."""
    dist = _dist()
    _, world = world_info(group)
    box = [header]
    if world > 1:
        dist.broadcast_object_list(box, src=src, group=group)
    return box[0]


class ShardedPhotometry(object):
    """One exposure replicated on every rank's GPU, star lists sharded (one process per GPU).

    Only ``src`` needs to hold pixel data (numpy arrays or tensors); the other ranks pass ``None`` and learn shape,
    dtypes and which of noise / mask exist from a header that ``src`` broadcasts first (a float64 image, or a plain
    ``pkfit`` set-up without noise and mask, therefore cannot desynchronise the NCCL broadcasts):

        sp = ShardedPhotometry(image, gauss, psf, ronois, phpadu, noise, mask)
        a = sp.aper(xc, yc, apr=np.array([3., 5., 8.]), skyrad=[40, 50], exact=True, ...)
        f = sp.fit('pkfit_noise', scale, x, y, a['sky'], radius=5)

    ``xc, yc ...`` must be identical on every rank (use ``broadcast_columns`` when only one rank has
    them).  Results are dicts of CUDA tensors in the caller's star order, identical on every rank.
    ``shape`` / ``image_dtype`` / ``with_noise`` are accepted for backward compatibility and checked against the header."""

    def __init__(self, image, gauss, psf, ronois, phpadu, noiseim=None, maskim=None, shape=None,
                 image_dtype=None, src=0, group=None, device=None, with_noise=None):
        from . import _lib
        from ._pkfit_core import PkfitEngine
        torch = _lib.torch_cuda()
        self.group, self.src = group, src
        self.rank, self.world = world_info(group)
        dev = _lib.device_of(device)
        self.device = dev
        img_t = noise_t = mask_t = None
        if self.rank == src:
            img_t = _lib.to_device_image(image, dev)
            noise_t = _lib.to_device_image(noiseim, dev) if noiseim is not None else None
            mask_t = _lib.to_device_mask(maskim, dev) if maskim is not None else None
        hdr = broadcast_header(exposure_header(img_t, noise_t, mask_t) if self.rank == src else None, src=src, group=group)
        if shape is not None and tuple(shape) != tuple(hdr["shape"]):
            raise ValueError("shape=%r but the source rank holds a %r image" % (tuple(shape), tuple(hdr["shape"])))
        if with_noise is not None and bool(with_noise) != (hdr["noise_dtype"] is not None):
            raise ValueError("with_noise=%r disagrees with the source rank" % (with_noise,))
        if self.rank != src:
            dt = {"torch.float32": torch.float32, "torch.float64": torch.float64}
            img_t = torch.empty(hdr["shape"], dtype=dt[hdr["image_dtype"]], device=dev)
            noise_t = torch.empty(hdr["shape"], dtype=dt[hdr["noise_dtype"]], device=dev) if hdr["noise_dtype"] else None
            mask_t = torch.empty(hdr["shape"], dtype=torch.uint8, device=dev) if hdr["has_mask"] else None
        broadcast_exposure([img_t, noise_t, mask_t], src=src, group=group)
        self.image_t = img_t
        self.engine = PkfitEngine(img_t, gauss, psf, ronois, phpadu, noise_t, mask_t, device=dev)
        self._side = torch.cuda.Stream(device=dev)

    def plan(self, x, y):
        xh = x.cpu().numpy() if hasattr(x, "cpu") else np.asarray(x)
        yh = y.cpu().numpy() if hasattr(y, "cpu") else np.asarray(y)
        return ShardPlan.for_stars(xh, yh, self.image_t.shape[0], self.image_t.shape[1], self.world)

    def aper_async(self, xc, yc, plan=None, **kw):
        """Sharded ``aper.aper`` (aper.py:26-433), keywords as ``aper.aper_device``.  Returns ``(shard_results, handle)``:
        this rank's own result columns right away (device tensors, e.g. the sky a following fit needs) and a
        ``GatherHandle`` whose all-gather + un-permute run on a side stream, hidden by whatever is launched next."""
        from . import _lib
        from .aper import aper_device
        torch = _lib.torch_cuda()
        plan = plan or self.plan(xc, yc)
        x_s = plan.take(_lib.to_device_f64(xc, self.device), self.rank)
        y_s = plan.take(_lib.to_device_f64(yc, self.device), self.rank)
        naper = _lib.host_f64(kw.get("apr", 5)).size
        arena = ResultArena(APER_SPEC(torch, naper), plan.m, self.device)
        aper_device(self.image_t, x_s, y_s, out=arena.views, **kw)
        return arena.views, gather_arena(plan, arena, group=self.group, side_stream=self._side)

    def aper(self, xc, yc, plan=None, **kw):
        return self.aper_async(xc, yc, plan=plan, **kw)[1].result()

    def fit_async(self, method, scale, x, y, sky, radius, maxiter=25, recenter=True, plan=None, sharded_inputs=None):
        """Sharded ``pkfit`` / ``pkfit_noise`` / ``pkfit_norecent_noise``.  ``sharded_inputs`` = dict with this rank's shard
        of any of scale / x / y / sky already on the device (e.g. ``aper_async``'s own ``sky`` column), which then are
        not taken from the full-length arguments."""
        from . import _lib
        torch = _lib.torch_cuda()
        variant = {"pkfit": _lib.PKFIT, "pkfit_noise": _lib.PKFIT_NOISE, "pkfit_norecent_noise": _lib.PKFIT_NORECENT}[method]
        plan = plan or self.plan(x, y)
        n = plan.n
        cols = {}
        for name, full in (("scale", scale), ("x", x), ("y", y), ("sky", sky)):
            if sharded_inputs and name in sharded_inputs:
                cols[name] = sharded_inputs[name]
            else:
                cols[name] = plan.take(_lib.to_device_f64(full, self.device, n), self.rank)
        arena = ResultArena(FIT_SPEC(torch), plan.m, self.device)
        self.engine.fit_device(variant, cols["scale"], cols["x"], cols["y"], cols["sky"], radius, maxiter=maxiter,
                               recenter=recenter, out=arena.views)
        return arena.views, gather_arena(plan, arena, group=self.group, side_stream=self._side)

    def fit(self, method, scale, x, y, sky, radius, maxiter=25, recenter=True, plan=None):
        return self.fit_async(method, scale, x, y, sky, radius, maxiter=maxiter, recenter=recenter, plan=plan)[1].result()
