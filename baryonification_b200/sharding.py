"""Multi-GPU sharding of the runners: one process per GPU, ``torch.distributed`` for the plumbing.

SURVEY.md section 8(e).  The reference's only parallelism is joblib over halo sub-catalogues
(``utils/Parallelize.py``: ``SimpleParallel`` :92-113, ``SplitJoinParallel`` :213-318, which works for the
*linear* Paint runners only, :209).  Here the OUTPUT is partitioned instead:

* shell runners: contiguous HEALPix RING pixel ranges made of whole 8-ring bands with (nearly) equal pixel
  counts; every rank receives the halos whose query disc can touch its bands (catalogue order preserved, so the
  per-pixel summation order -- and the bits -- equal the single-GPU run).  PaintProfilesShell needs no reduction
  at all (pixel owners); BaryonifyShell re-bins its own pixels into a full-sphere accumulator, and the deposits
  that land outside the rank's range (the boundary mass, a few rings) are sent to their owners and added there in
  rank order -- the only data-path collective;
* snapshot runner: particles are split by INDEX (any partition works, the kernel is particle-centric), the halo
  grid is replicated (it is ~10 MB); no collective on the data path, output order = input order;
* grid runners: halos are split across ranks, the per-cell accumulators are summed with one all-reduce.

Everything in this file is device-agnostic torch + numpy, so the N > 1 logic is covered on CPU with the gloo
backend (tests/test_sharding.py); the kernels run on each rank's own device through the engines.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from . import _lib


# ----------------------------------------------------------------------------------------------------------
# partitions
# ----------------------------------------------------------------------------------------------------------
def band_pixel_range(nside, band_lo, band_hi):
    lib = _lib.load()
    f, l = C.c_int64(), C.c_int64()
    _lib.check(lib.bf_shell_band_pixel_range(nside, band_lo, band_hi, C.byref(f), C.byref(l)), "bf_shell_band_pixel_range")
    return f.value, l.value


_band_ends_cache = {}


def band_ends(nside):
    """ends[b] = first pixel after ring bands 0 .. b - 1 (int64 array of bf_shell_num_bands + 1 entries; cached per NSIDE:
    a few thousand library calls, milliseconds that every ``process()`` would otherwise pay again)."""
    nside = int(nside)
    ends = _band_ends_cache.get(nside)
    if ends is None:
        nb = int(_lib.load().bf_shell_num_bands(nside))
        ends = np.array([band_pixel_range(nside, 0, b)[1] for b in range(nb + 1)], dtype=np.int64)
        ends.flags.writeable = False
        _band_ends_cache[nside] = ends
    return ends


def band_split(nside, world):
    """``world`` contiguous ring-band ranges [lo, hi) with (nearly) equal pixel counts, covering the sphere."""
    lib = _lib.load()
    nb = int(lib.bf_shell_num_bands(nside))
    if world > nb:
        raise ValueError(f"cannot split {nb} ring bands of NSIDE {nside} over {world} ranks")
    npix = 12 * nside * nside
    ends = band_ends(nside)
    cuts = [0]
    for r in range(1, world):
        c = int(np.searchsorted(ends, npix * r / world))
        cuts.append(min(max(c, cuts[-1] + 1), nb - (world - r)))
    cuts.append(nb)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def ring_colatitude(nside, ring):
    """theta of ring ``ring`` (1 .. 4 nside - 1); 0 and pi for the virtual rings 0 and 4 nside."""
    ring = np.asarray(ring, dtype=np.float64)
    n = float(nside)
    z = np.where(ring < n, 1.0 - ring * ring / (3.0 * n * n),
                 np.where(ring <= 3 * n, (2.0 * n - ring) * 2.0 / (3.0 * n), -1.0 + (4 * n - ring) ** 2 / (3.0 * n * n)))
    return np.arccos(np.clip(z, -1.0, 1.0))


def halos_touching_bands(vec_z, radius, nside, band_lo, band_hi, margin_rings=3):
    """Mask of the halos whose query disc (plus a safety margin of ``margin_rings`` rings, which also covers the
    4-neighbour fallback of tiny discs) can touch ring bands [band_lo, band_hi).  Conservative: the binning kernel
    makes the exact per-band decision."""
    theta0 = np.arccos(np.clip(np.asarray(vec_z, dtype=np.float64), -1.0, 1.0))
    r = np.asarray(radius, dtype=np.float64)
    first_ring = max(band_lo * 8 + 1 - margin_rings, 0)
    last_ring = min(band_hi * 8 + margin_rings, 4 * nside)
    t_lo, t_hi = ring_colatitude(nside, first_ring), ring_colatitude(nside, last_ring)
    keep = (theta0 + r >= t_lo) & (theta0 - r <= t_hi)
    return keep | ~np.isfinite(r)


def index_split(n, world):
    """``world`` contiguous index ranges of (nearly) equal length covering [0, n)."""
    edges = [(n * r) // world for r in range(world + 1)]
    return [(edges[r], edges[r + 1]) for r in range(world)]


# ----------------------------------------------------------------------------------------------------------
# collectives (any backend / device)
# ----------------------------------------------------------------------------------------------------------
def _world(group):
    if not dist.is_initialized():  # a single process without a process group is the 1-rank case
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def _touched_extent(acc, first, last, window=None):
    """``[lo, hi, widen]`` (int64 tensor on acc's device, NO host synchronisation): [lo, hi) is the extent of the non-zero
    entries of ``acc`` outside the own range [first, last) (lo = first / hi = last if none).  ``window``: look at that
    many entries next to the range only (displaced mass lands a few rings away); ``widen`` = 1 when the window's far
    edge is itself non-zero, i.e. the whole array has to be scanned (``window=None``) to be sure."""
    n = acc.numel()
    dev = acc.device
    lo = torch.full((), first, dtype=torch.int64, device=dev)
    hi = torch.full((), last, dtype=torch.int64, device=dev)
    widen = torch.zeros((), dtype=torch.bool, device=dev)
    flat = acc.reshape(n, -1) if acc.dim() > 1 else acc
    a = 0 if window is None else max(0, first - window)
    if first > a:
        nz = flat[a:first] != 0
        nz = nz.any(dim=1) if nz.dim() > 1 else nz
        any_ = nz.any()
        i0 = torch.argmax(nz.to(torch.uint8))  # index of the FIRST non-zero entry
        lo = torch.where(any_, a + i0, lo)
        if a > 0:
            widen = widen | (any_ & (i0 == 0))
    b = n if window is None else min(n, last + window)
    if last < b:
        nz = flat[last:b] != 0
        nz = nz.any(dim=1) if nz.dim() > 1 else nz
        any_ = nz.any()
        i1 = (b - last - 1) - torch.argmax(nz.flip(0).to(torch.uint8))  # index of the LAST non-zero entry
        hi = torch.where(any_, last + i1 + 1, hi)
        if b < n:
            widen = widen | (any_ & (i1 == b - last - 1))
    return torch.stack([lo, hi, widen.to(torch.int64)])


def exchange_boundary(acc, ranges, group=None, window=None):
    """Boundary mass exchange of the re-binning stage.

    ``acc`` is this rank's full-length accumulator: its own deposits, mostly inside its own range
    ``ranges[rank] = [first, last)`` plus the few that landed outside (displaced mass crossing the shard
    boundary).  Every rank sends those outside parts to their owners; the owner adds what it receives to its own
    slice in SOURCE-RANK order (a fixed order: bit-reproducible).  Returns the finished own slice (a view of acc).
    The touched extents are found on the device and reach the host with ONE read-back (after the all-gather); a rank
    whose window turned out too narrow rescans its whole array and the extents are gathered once more.
    """
    rank, world = _world(group)
    first, last = ranges[rank]
    own = acc[first:last]
    if world == 1:
        return own

    def gathered(ext):
        parts = [torch.empty_like(ext) for _ in range(world)]
        dist.all_gather(parts, ext, group=group)
        return torch.stack(parts).cpu().tolist()

    all_ext = gathered(_touched_extent(acc, first, last, window))
    if any(e[2] for e in all_ext):  # (every rank sees the same flags, so every rank takes this branch)
        mine = _touched_extent(acc, first, last, None) if all_ext[rank][2] else \
            torch.tensor(all_ext[rank], dtype=torch.int64, device=acc.device)
        all_ext = gathered(mine)
    ops, recvs = [], []
    for src in range(world):
        s_first, s_last = ranges[src]
        s_lo, s_hi = all_ext[src][0], all_ext[src][1]
        for dst in range(world):
            if dst == src:
                continue
            d_first, d_last = ranges[dst]
            for a, b in ((max(s_lo, d_first), min(s_first, d_last)), (max(s_last, d_first), min(s_hi, d_last))):
                if b <= a:
                    continue
                if rank == src:
                    ops.append(dist.P2POp(dist.isend, acc[a:b].contiguous(), dist.get_global_rank(group, dst)
                                          if group is not None else dst, group=group))
                elif rank == dst:
                    buf = torch.empty(b - a, dtype=acc.dtype, device=acc.device)
                    ops.append(dist.P2POp(dist.irecv, buf, dist.get_global_rank(group, src)
                                          if group is not None else src, group=group))
                    recvs.append((src, a, b, buf))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    for src, a, b, buf in sorted(recvs, key=lambda t: (t[0], t[1])):
        acc[a:b] += buf
    return own


def gather_slices(own, ranges, group=None, out=None):
    """All ranks' slices -> the full array on every rank.  Slices have different lengths (whole ring bands), so this is
    an all-gather-v: one batched group of point-to-point copies in which every rank sends its slice to every peer and
    receives the peers' slices IN PLACE at ``out[a:b]`` -- no padding, no staging buffers, the only bytes moved are
    the ones that cross the links.  ``own`` may already be ``out[ranges[rank]]`` (then nothing is copied locally)."""
    rank, world = _world(group)
    n = ranges[-1][1]
    if world == 1:
        return own
    tail = own.shape[1:]
    if out is None:
        out = torch.empty((n,) + tuple(tail), dtype=own.dtype, device=own.device)
    a, b = ranges[rank]
    mine = out[a:b]
    if b > a and mine.data_ptr() != own.data_ptr():
        mine.copy_(own)
    ops = []
    peer = (lambda r: dist.get_global_rank(group, r)) if group is not None else (lambda r: r)
    for r, (ra, rb) in enumerate(ranges):
        if r == rank:
            continue
        if b > a:
            ops.append(dist.P2POp(dist.isend, mine, peer(r), group=group))
        if rb > ra:
            ops.append(dist.P2POp(dist.irecv, out[ra:rb], peer(r), group=group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return out


MARGIN_RINGS = 96  # = BF_SHELL_GATHER_MARGIN_RINGS / 2: the farthest the shard form of the destination gather deposits


def margin_pixels(nside):
    return MARGIN_RINGS * 4 * int(nside)


def fixed_margins_ok(ranges, F):
    """The fixed-width neighbour exchange applies when every rank's range is at least one margin long, so that what a
    rank deposits beyond its range lands on the ADJACENT rank only.  Decided from the partition alone: every rank
    takes the same branch."""
    return len(ranges) > 1 and all(b - a >= F for a, b in ranges)


def exchange_margins(local, ranges, F, group=None):
    """Boundary mass exchange with FIXED margins: ``local`` holds this rank's partial sums over
    [first - F, last + F) (without the side that has no neighbour); the two margins go to the adjacent ranks, the
    margins received from them are added to the first / last F own pixels (each boundary region has exactly one
    foreign contributor, so the order is fixed).  No size negotiation, no host read-back: four point-to-point messages
    of F pixels.  Returns the finished own slice (a view of ``local``)."""
    rank, world = _world(group)
    first, last = ranges[rank]
    lo_n, hi_n = rank > 0, rank < world - 1
    off0 = F if lo_n else 0
    own = local[off0:off0 + (last - first)]
    if world == 1:
        return own
    peer = (lambda r: dist.get_global_rank(group, r)) if group is not None else (lambda r: r)
    ops, r_lo, r_hi = [], None, None
    if lo_n:
        r_lo = torch.empty(F, dtype=local.dtype, device=local.device)
        ops += [dist.P2POp(dist.isend, local[:F], peer(rank - 1), group=group),
                dist.P2POp(dist.irecv, r_lo, peer(rank - 1), group=group)]
    if hi_n:
        r_hi = torch.empty(F, dtype=local.dtype, device=local.device)
        ops += [dist.P2POp(dist.isend, local[local.numel() - F:], peer(rank + 1), group=group),
                dist.P2POp(dist.irecv, r_hi, peer(rank + 1), group=group)]
    for w in dist.batch_isend_irecv(ops):
        w.wait()
    if r_lo is not None:
        own[:F] += r_lo
    if r_hi is not None:
        own[own.numel() - F:] += r_hi
    return own


# ----------------------------------------------------------------------------------------------------------
# grid slabs
# ----------------------------------------------------------------------------------------------------------
def halos_touching_rows(x, bins, Nsize, a_first, a_last):
    """Mask of the halos whose cut-out box (Map2DRunner.py:494-512: ``Nsize`` cells around ``argmin|bins - x|``,
    periodic) can reach rows ``a_first <= a < a_last`` of the first grid axis.  One extra row either side stands in
    for the exact argmin, so the mask is a superset; the kernel's membership test is the exact one."""
    n = bins.size
    res = bins[1] - bins[0]
    cen = np.clip(np.rint((np.asarray(x, dtype=np.float64) - bins[0]) / res), 0, n - 1).astype(np.int64)
    W = np.asarray(Nsize, dtype=np.int64)
    lo, length = cen - W // 2 - 1, W + 2
    keep = length >= n
    for shift in (-n, 0, n):
        keep |= (lo + shift < a_last) & (lo + shift + length > a_first)
    return keep & (W >= 2)


def ghost_rows(rows, rank, below, above, n):
    """Global row indices (periodic) that ``rank`` needs beyond its own slab: ``below`` rows before it and ``above``
    rows after it, without the rows it owns itself."""
    a0, a1 = rows[rank]
    need = [(a0 - k) % n for k in range(min(below, n), 0, -1)] + [(a1 + k) % n for k in range(min(above, n))]
    seen, out = set(), []
    for g in need:
        if not (a0 <= g < a1) and g not in seen:
            seen.add(g)
            out.append(g)
    return out


def exchange_rows(buf, rows, below, above, group=None):
    """Fill the ghost rows of ``buf`` ([n_rows, ...] with GLOBAL row indexing; every rank holds valid data in its
    own slab ``rows[rank]``) from their owners.  Point-to-point only, one message per (owner, needer) pair."""
    rank, world = _world(group)
    n = buf.shape[0]
    if world == 1:
        return buf
    owner = np.empty(n, dtype=np.int64)
    for r, (a, b) in enumerate(rows):
        owner[a:b] = r
    ops, recvs = [], []
    for needer in range(world):
        need = ghost_rows(rows, needer, below, above, n)
        for src in range(world):
            idx = [g for g in need if owner[g] == src]
            if not idx or src == needer:
                continue
            if rank == src:
                payload = buf[torch.as_tensor(idx, device=buf.device)].contiguous()
                ops.append(dist.P2POp(dist.isend, payload, needer, group=group))
            elif rank == needer:
                box = torch.empty((len(idx),) + tuple(buf.shape[1:]), dtype=buf.dtype, device=buf.device)
                ops.append(dist.P2POp(dist.irecv, box, src, group=group))
                recvs.append((idx, box))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    for idx, box in recvs:
        buf[torch.as_tensor(idx, device=buf.device)] = box
    return buf


# ----------------------------------------------------------------------------------------------------------
# sharded runners (call from every rank of an initialised process group with the SAME host inputs)
# ----------------------------------------------------------------------------------------------------------
def _subset(scalars, keep):
    return {k: (np.ascontiguousarray(v[keep]) if isinstance(v, np.ndarray) and v.shape[:1] == keep.shape else v)
            for k, v in scalars.items() if k != "_pinned"}  # (the page-locked slabs hold the WHOLE catalogue)


class SharedHostArena:
    """A host array that every rank of a one-box job maps (POSIX shared memory under /dev/shm) and page-locks
    (``cudaHostRegister``), so that each rank's GPU writes ITS piece of a result straight into the caller-visible array
    over its own PCIe link -- N links in parallel instead of one rank pulling the whole map through one.  Rank 0 names
    and creates the segment, the others attach after a barrier; the file is unlinked as soon as everybody has it mapped
    (the memory lives as long as the mappings).  Without CUDA (the gloo tests) the array is simply shared."""

    def __init__(self, n, group=None, dtype=np.float64):
        import os
        import uuid
        rank, world = _world(group)
        self.n, self.registered, self._tensor = int(n), False, None
        nbytes = max(self.n, 1) * np.dtype(dtype).itemsize
        name = [None]
        if rank == 0:
            name[0] = "/dev/shm/bf200_%d_%s" % (os.getpid(), uuid.uuid4().hex[:12])
            with open(name[0], "wb") as f:
                f.truncate(nbytes)
        if world > 1:
            dist.broadcast_object_list(name, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        self.array = np.memmap(name[0], dtype=dtype, mode="r+", shape=(max(self.n, 1),))[:self.n]
        if world > 1:
            dist.barrier(group=group)
        if rank == 0:
            os.unlink(name[0])
        if torch.cuda.is_available():
            import ctypes as C
            t = torch.from_numpy(self.array)
            rc = int(_lib.load(require_cuda=True).bf_host_pin(C.c_void_p(t.data_ptr()), t.numel() * t.element_size())) \
                if t.numel() else -1
            self.registered = rc == 0
            self._tensor = t
        self.all_registered = self.registered  # page-locked on EVERY rank: what collective decisions must be based on
        if world > 1:
            flag = torch.tensor([1 if self.registered else 0], dtype=torch.int32,
                                device="cuda" if dist.get_backend(group) == "nccl" else "cpu")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
            self.all_registered = bool(flag.item())

    @property
    def tensor(self):
        if self._tensor is None:
            self._tensor = torch.from_numpy(self.array)
        return self._tensor

    def close(self):
        if self.registered:
            import ctypes as C
            _lib.load().bf_host_unpin(C.c_void_p(self._tensor.data_ptr()))
            self.registered = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def claim_host_threads():
    """torchrun starts every rank with OMP_NUM_THREADS=1, which makes torch's pageable -> page-locked staging copies (the
    map slice, the particle records: GBs per call) single-threaded.  The sharded runners give torch this rank's share of
    the host cores instead (never fewer threads than it already has)."""
    from ._device import thread_budget
    return thread_budget()


_SHELL_KEYS = ("a", "D", "radius", "z_in", "M_in", "lon", "lat", "cut", "lnr_shift")


class ShardedShell:
    """Ring-band sharded BaryonifyShell / PaintProfilesShell (one process per GPU of one box).

    ``runner`` is an (unprocessed) ``BaryonifyShell`` or ``PaintProfilesShell`` built identically on every rank.
    ``process()`` returns the full map on every rank, like the single-GPU runner -- as a view of a host array all ranks
    share (``SharedHostArena``; two of them alternate, so a result stays valid until the second-next call):

    * host set-up: every rank prepares the per-halo scalars of 1/N of the catalogue, one device all-gather (~100 B per
      halo over NVLink) hands every rank the whole table, and each keeps the halos whose footprint can touch its bands
      (catalogue order preserved: the per-pixel summation order -- and the bits -- equal the single-GPU run);
    * pair stage on the rank's bands: no collective (pixel owners);
    * BaryonifyShell re-binning: the rank's partial sums over its range widened by a fixed margin, the two margins go
      to the adjacent ranks and are added there (``exchange_margins``: four messages, no negotiation) -- the boundary
      mass exchange, the only data-path collective; small spheres whose bands are shorter than the margin take the
      general ``exchange_boundary``;
    * every rank copies ITS finished band to the shared host array over its own PCIe link."""

    def __init__(self, runner, group=None, device=None):
        from .Runners import shell as _shell
        self.runner, self.group = runner, group
        self.baryonify = isinstance(runner, _shell.BaryonifyShell)
        self.rank, self.world = _world(group)
        self.device = device
        self.npairs = None
        self.timings = {}
        self._arenas, self._turn = [None, None], 0
        claim_host_threads()

    def plan(self):
        nside = self.runner.LightconeShell.NSIDE
        bands = band_split(nside, self.world)
        ranges = [band_pixel_range(nside, lo, hi) for lo, hi in bands]
        return bands, ranges

    # ---- per-halo scalars: 1/N of the host work per rank, all-gathered on the device ------------------------
    def scalars_device(self, dev, bands):
        """dict of device tensors (the ``ShellEngine`` inputs) for the halos that can touch ``bands``."""
        import copy
        from . import host
        r = self.runner
        cat = r.HaloLightConeCatalog
        n = len(cat.cat)
        a, b = index_split(n, self.world)[self.rank]
        part = copy.copy(cat)
        part.cat = cat.cat[a:b]
        import time
        t_host = time.perf_counter()
        s = host.shell_scalars(part, r.cosmo, r.mass_def, r.epsilon_max, r.model, self.baryonify)
        self.timings["host_scalars"] = time.perf_counter() - t_host  # (part of host_setup_allgather)
        ex = host.property_columns(part.cat, r.model)
        nx = 0 if ex is None else ex.shape[1]
        nk = len(_SHELL_KEYS)
        width = nk + 3 + nx
        chunk = -(-n // self.world) if n else 0  # equal pieces for one all_gather_into_tensor
        # one ROW per scalar (key-major): the page-locked slabs the host set-up wrote go up as they are, one asynchronous
        # copy each -- no host-side interleave into a [halo][key] staging array
        pins = s.get("_pinned") or {}
        mine_d = torch.zeros((width, chunk), dtype=torch.float64, device=dev)

        def up(k):
            src = pins[k] if k in pins else torch.from_numpy(np.ascontiguousarray(s[k], dtype=np.float64))
            return src.to(dev, non_blocking=True)

        for c, k in enumerate(_SHELL_KEYS):
            mine_d[c, :b - a].copy_(up(k), non_blocking=True)
        mine_d[nk:nk + 3, :b - a] = up("vec").t()
        if nx:
            mine_d[nk + 3:, :b - a] = torch.from_numpy(ex).to(dev).t()
        if self.world > 1:
            full = torch.empty((self.world, width, chunk), dtype=torch.float64, device=dev)
            dist.all_gather_into_tensor(full.view(-1), mine_d.view(-1), group=self.group)
            sizes = [hi - lo for lo, hi in index_split(n, self.world)]
            if all(sz == chunk for sz in sizes):
                full = full.permute(1, 0, 2).reshape(width, chunk * self.world)
            else:  # drop the padding columns of the shorter pieces
                full = torch.cat([full[i, :, :sz] for i, sz in enumerate(sizes)], dim=1)
        else:
            full = mine_d[:, :n]
        nside = r.LightconeShell.NSIDE
        ci = {k: c for c, k in enumerate(_SHELL_KEYS)}
        if self.world > 1:  # halos_touching_bands on the device (conservative by `margin_rings`, like the host form)
            lo, hi = bands
            theta0 = torch.acos(full[nk + 2].clamp(-1.0, 1.0))
            rad = full[ci["radius"]]
            first_ring, last_ring = max(lo * 8 + 1 - 3, 0), min(hi * 8 + 3, 4 * nside)
            t_lo, t_hi = float(ring_colatitude(nside, first_ring)), float(ring_colatitude(nside, last_ring))
            keep = ((theta0 + rad >= t_lo) & (theta0 - rad <= t_hi)) | ~torch.isfinite(rad)
            full = full[:, keep]
        t = {k: full[ci[k]].contiguous() for k in _SHELL_KEYS}
        t["vec"] = full[nk:nk + 3].t().contiguous()
        t["extras"] = full[nk + 3:].t().contiguous() if nx else None
        return t

    def process_device(self, gather=False, out_h=None):
        """Device part: returns (own slice [first, last) on the device -- or the whole map with ``gather=True`` --,
        (sum new, sum old) or None, ranges)."""
        import time
        from . import _device as dv, engine
        from .tables import displacement_host_table, profile_host_table
        r = self.runner
        r._check_keys(self.baryonify)
        dev = dv.device(self.device)
        nside = r.LightconeShell.NSIDE
        bands, ranges = self.plan()
        lo, hi = bands[self.rank]
        first, last = ranges[self.rank]
        mode = engine.BARYONIFY if self.baryonify else engine.PAINT
        htab = displacement_host_table(r.model) if self.baryonify else profile_host_table(r.model, "2D")
        with torch.cuda.device(dev):
            t0 = time.perf_counter()
            t = self.scalars_device(dev, (lo, hi))
            t0 = self._tick("host_setup_allgather", t0, dev)
            up = None

            def start_upload():  # this rank's slice of the map travels under the pair kernel
                orig = np.ascontiguousarray(r.LightconeShell.map, dtype=np.float64).reshape(-1)
                return dv.ChunkedUpload(orig[first:last], dev, np.linspace(0, last - first, 9).astype(np.int64))

            early = self.baryonify and not dv.pinned_already(
                np.ascontiguousarray(r.LightconeShell.map, dtype=np.float64).reshape(-1)[first:last])
            if early:  # pageable: the staging memcpy starts now (after the host set-up, which wanted the cores); a slice
                up = start_upload()  # that is page-locked in place starts after the binning stage's small uploads
            eng = engine.ShellEngine(nside, mode, htab, None, dev, variant="gather", bands=(lo, hi))
            eng.max_entries = getattr(self, "_bin_hint", 0)  # (a repeated call skips the counting pass of the binning)
            eng.upload_device(t).bin()
            if self.baryonify and up is None:
                up = start_upload()
            if eng.nentries is not None:
                self._bin_hint = eng.nentries + eng.nentries // 16 + 1024
            t0 = self._tick("collapse_bin", t0, dev)
            if out_h is not None and not self.baryonify:  # paint: finished band groups travel under the next group's kernel
                eng.accumulate_to_host(out_h=out_h)
                pairs = eng.cnt.clone()
                if self.world > 1:
                    dist.all_reduce(pairs, group=self.group)
                self.npairs = int(pairs.item())
                self._tick("pairs_d2h", t0, dev)
                return None, None, ranges
            acc = eng.accumulate()
            pairs = eng.cnt.clone()
            if self.world > 1:
                dist.all_reduce(pairs, group=self.group)
            if not self.baryonify:
                self.npairs = int(pairs.item())
                self._tick("pairs", t0, dev)
                return (gather_slices(acc, ranges, self.group) if gather else acc), None, ranges
            t0 = self._tick("pairs", t0, dev)
            orig_d = up.result()
            F = margin_pixels(nside)
            if fixed_margins_ok(ranges, F):
                local, stray = eng.regrid_shard(orig_d, acc, F, self.rank > 0, self.rank < self.world - 1, method=r.regrid)
                own = exchange_margins(local, ranges, F, self.group)
            else:  # small spheres: bands shorter than the margin
                stray = None
                new_full = eng.regrid(orig_d, acc, method=r.regrid)
                own = exchange_boundary(new_full, ranges, self.group, window=64 * 4 * nside)
            sums = torch.stack([own.sum(), orig_d.sum(), (stray if stray is not None else own.new_zeros(())).to(own.dtype)])
            if self.world > 1:
                dist.all_reduce(sums, group=self.group)  # the reference's mass check, HealpixRunner.py:344-346
            self.npairs = int(pairs.item())
            sums = sums.tolist()
            self._tick("regrid_exchange", t0, dev)
            if sums[2] != 0:
                raise _lib.BFError("ShardedShell: displaced mass landed more than %d rings outside a shard" % MARGIN_RINGS)
            return (gather_slices(own, ranges, self.group) if gather else own), (sums[0], sums[1]), ranges

    def _tick(self, name, t0, dev):
        import time
        torch.cuda.current_stream(dev).synchronize()  # (this stream only: the map upload on its side stream runs on)
        self.timings[name] = time.perf_counter() - t0
        return time.perf_counter()

    def _arena(self, n):
        self._turn ^= 1
        ar = self._arenas[self._turn]
        if ar is None or ar.n != n:
            ar = self._arenas[self._turn] = SharedHostArena(n, self.group)
        return ar

    def process(self):
        import time
        if not self.baryonify and torch.cuda.is_available():  # paint: read-back pipelined with the pair stage
            _, ranges = self.plan()
            arena = self._arena(ranges[-1][1])
            a, b = ranges[self.rank]
            if arena.all_registered:
                self.process_device(gather=False, out_h=arena.tensor[a:b])
                t0 = time.perf_counter()
                if self.world > 1:
                    dist.barrier(group=self.group)
                self.timings["barrier"] = time.perf_counter() - t0
                return arena.array
            self._turn ^= 1  # (not page-locked: the plain path below takes the same arena)
        out, sums, ranges = self.process_device(gather=False)
        t0 = time.perf_counter()
        if sums is not None:
            assert np.isclose(sums[0], sums[1]), \
                "ERROR in pixel regridding, sum(new_map) [%0.14e] != sum(oldmap) [%0.14e]" % sums
        arena = self._arena(ranges[-1][1])
        a, b = ranges[self.rank]
        t1 = time.perf_counter()
        arena.tensor[a:b].copy_(out, non_blocking=arena.registered)  # every rank: its band, its own PCIe link
        if out.is_cuda:
            torch.cuda.current_stream(out.device).synchronize()
        t2 = time.perf_counter()
        if self.world > 1:
            dist.barrier(group=self.group)
        self.timings["arena"], self.timings["d2h"], self.timings["barrier"] = t1 - t0, t2 - t1, time.perf_counter() - t2
        self.timings["d2h_barrier"] = time.perf_counter() - t0
        return arena.array


class ShardedSnapshot:
    """Index-sharded BaryonifySnapshot: every rank displaces a contiguous slice of the particle array against the
    full (replicated) halo grid.  No collective on the data path.  ``process()`` returns the whole displaced array on
    every rank: for the reference's (M, x, y, z) record layout as a view of a host array all ranks of the box share
    (``SharedHostArena``; every rank's GPU writes ITS slice over its own PCIe link, two arenas alternate so a result
    stays valid until the second-next call), for any other layout by an object all-gather."""

    def __init__(self, runner, group=None, device=None):
        self.runner, self.group, self.device = runner, group, device
        self.rank, self.world = _world(group)
        self.npairs = None
        self._arenas, self._turn = [None, None], 0
        claim_host_threads()

    def _arena(self, n):
        self._turn ^= 1
        ar = self._arenas[self._turn]
        if ar is None or ar.n != n:
            ar = self._arenas[self._turn] = SharedHostArena(n, self.group)
        return ar

    def process(self, gather=True):
        import copy
        from .Runners.snapshot import BaryonifySnapshot
        r = self.runner
        cat = r.ParticleSnapshot.cat
        ranges = index_split(len(cat), self.world)
        a, b = ranges[self.rank]
        part = copy.copy(r.ParticleSnapshot)
        part.cat = cat[a:b]
        local = type(r)(r.HaloNDCatalog, part, r.epsilon_max, r.model, r.mass_def, verbose=False,
                        device=self.device)
        shared = None
        if gather and torch.cuda.is_available() and BaryonifySnapshot._is_reference_layout(cat) and len(cat) > 0:
            arena = self._arena(4 * len(cat))
            if arena.all_registered:
                shared = arena
            else:
                self._turn ^= 1
        if shared is not None:
            local.process(_out=shared.tensor.view(-1, 4)[a:b])
        else:
            out = local.process()
        pairs = torch.tensor([local.npairs], dtype=torch.int64,
                             device="cuda" if dist.get_backend(self.group) == "nccl" else "cpu")
        dist.all_reduce(pairs, group=self.group)  # (also: every rank's slice has arrived in the shared array)
        self.npairs = int(pairs.item())
        if shared is not None:
            return np.asarray(shared.array).view(cat.dtype).reshape(cat.shape)
        if not gather:
            return out
        parts = [None] * self.world
        dist.all_gather_object(parts, out, group=self.group)
        return np.concatenate(parts)


class ShardedGrid:
    """Slab-sharded BaryonifyGrid / PaintProfilesGrid: every rank owns a run of rows of the first grid axis.

    Pair loop: the rank evaluates only the cells of its slab (cell-owner gather, ``bf_grid_gather_rows``) against the
    halos whose cut-out can reach it -- catalogue order is kept inside the subset, so a cell's sum is bit-identical to
    the single-GPU one, and there is no collective.  Re-binning (BaryonifyGrid): the destination gather needs the
    offsets and masses of a few source rows either side of the slab; those INPUT ghost rows are exchanged point to
    point (``exchange_rows``) after the displacement window has been agreed on with a MIN / MAX all-reduce of six
    integers, and every rank then writes its own destination rows -- again the single-GPU bits.  Arrays keep global
    indexing (full-size allocations, slab-sized work).  ``process()`` returns the full map on every rank."""

    def __init__(self, runner, group=None, device=None):
        from .Runners import grid as _grid
        self.runner, self.group, self.device = runner, group, device
        self.baryonify = isinstance(runner, _grid.BaryonifyGrid)
        self.rank, self.world = _world(group)
        self.npairs = None
        claim_host_threads()

    def _local_runner(self, a0, a1):
        import copy
        from . import host
        r = self.runner
        G = r.GriddedMap
        s = host.grid_scalars(r.HaloNDCatalog, G, r.mass_def, r.epsilon_max, r.model, self.baryonify)
        keep = halos_touching_rows(s["pos"][:, 0], np.asarray(G.bins, dtype=np.float64), s["Nsize"], a0, a1)
        cat = copy.copy(r.HaloNDCatalog)
        cat.cat = r.HaloNDCatalog.cat[keep]
        local = copy.copy(r)
        local.HaloNDCatalog, local.rows, local.device, local.timings = cat, (a0, a1), self.device, {}
        return local

    GHOST_ROWS = 4  # rows of the map uploaded either side of the slab before the displacement window is known

    def upload(self, dev, ghost=None):
        """Everything the device stages need, resident in HBM: the halos that can reach this rank's slab (set-up on the
        host, table, collapsed rows), the slab's rows of the map plus ``ghost`` rows either side (more are fetched
        when the displacement window turns out wider).  Returns the state dict the stage methods take."""
        from . import _device as dv
        from .tables import displacement_host_table, profile_host_table
        r = self.runner
        G = r.GriddedMap
        n, dim = G.Npix, (2 if G.is2D else 3)
        rowlen = n ** (dim - 1)
        rows = index_split(n, self.world)
        a0, a1 = rows[self.rank]
        local = self._local_runner(a0, a1)
        ghost = self.GHOST_ROWS if ghost is None else int(ghost)
        with torch.cuda.device(dev):
            htab = displacement_host_table(r.model) if self.baryonify else profile_host_table(r.model, "2D" if G.is2D else "3D")
            table, s, desc, bins, keep = local._prepare(htab, self.baryonify, dev)
            st = dict(local=local, table=table, s=s, desc=desc, bins=bins, keep=keep, rows=rows, a0=a0, a1=a1, n=n, dim=dim,
                      rowlen=rowlen, dev=dev, cnt=dv.pair_counter(dev), have=set())
            if self.baryonify:
                st["orig_h"] = np.ascontiguousarray(G.map, dtype=np.float64).reshape(n, rowlen)
                st["orig"] = dv.zeros((n, rowlen), dev=dev) if self.world > 1 else None
                self._fetch_rows(st, list(range(a0, a1)) + ghost_rows(rows, self.rank, ghost, ghost, n))
                st["orig_slab"] = st["orig"][a0:a1]
                st["off"] = dv.zeros((n, rowlen * dim), dev=dev)
                st["out"] = dv.zeros((n, rowlen), dev=dev)
        return st

    def _fetch_rows(self, st, wanted):
        """Upload the rows of the map in ``wanted`` that are not on the device yet (global row indexing)."""
        from . import _device as dv
        new = sorted(set(int(g) for g in wanted) - st["have"])
        if st["orig"] is None:  # one rank: the whole map, one pinned copy
            st["orig"] = dv.to_dev(st["orig_h"], dev=st["dev"], pinned=True)
            st["have"] = set(range(st["n"]))
            return
        for g0, g1 in _runs(new):
            st["orig"][g0:g1] = dv.to_dev(st["orig_h"][g0:g1], dev=st["dev"], pinned=True)
        st["have"].update(new)

    def pairs_device(self, st):
        """The per-halo loop on this rank's slab (cell-owner gather, no collective): pixel offsets [n, rowlen * dim]
        (global row indexing, only the slab's rows are written) or the painted rows."""
        loc = st["local"]
        a0, a1 = st["a0"], st["a1"]
        out = st["off"] if self.baryonify else st.setdefault("paint", torch.zeros((st["n"], st["rowlen"]), dtype=torch.float64,
                                                                                   device=st["dev"]))
        out[a0:a1].zero_()
        st["cnt"].zero_()
        with torch.cuda.device(st["dev"]):
            from .Runners.grid import GRID_BARYONIFY, GRID_PAINT
            loc._accumulate(GRID_BARYONIFY if self.baryonify else GRID_PAINT, st["s"], st["desc"], st["bins"], st["table"],
                            out.view(-1, st["dim"]) if self.baryonify else out.view(-1), st["cnt"], st["dev"])
        self.local_pairs = loc.npairs
        return out

    def regrid_device(self, st, off):
        """BaryonifyGrid re-binning of the slab: agree on the displacement window (MIN / MAX all-reduce of six integers),
        exchange the ghost rows of the offsets, fetch missing ghost rows of the map, gather this rank's destination
        rows.  Returns the slab's rows of the new map (device, [a1 - a0, rowlen])."""
        import ctypes as C
        from . import _device as dv, _lib
        dev, n, dim, a0, a1, rows = st["dev"], st["n"], st["dim"], st["a0"], st["a1"], st["rows"]
        with torch.cuda.device(dev):
            lib = _lib.load(require_cuda=True)
            if "rg_ws" not in st:
                st["rg_ws"] = torch.empty(lib.bf_grid_regrid_gather_ws_bytes(), dtype=torch.uint8, device=dev)
            ws = st["rg_ws"]
            lohi = (C.c_int * 6)()
            _lib.call("bf_grid_regrid_window", dim, n, dv.ptr(off), a0, a1, dv.ptr(ws), ws.numel(), lohi, dv.stream_ptr())
            if self.world > 1:
                lo_t = torch.tensor(list(lohi)[:3], dtype=torch.int64, device=dev)
                hi_t = torch.tensor(list(lohi)[3:], dtype=torch.int64, device=dev)
                dist.all_reduce(lo_t, op=dist.ReduceOp.MIN, group=self.group)
                dist.all_reduce(hi_t, op=dist.ReduceOp.MAX, group=self.group)
                lohi = (C.c_int * 6)(*[int(v) for v in lo_t.tolist() + hi_t.tolist()])
            below, above = max(0, lohi[3] + 1), max(0, -lohi[0])  # source rows a0 - hi - 1 ... a1 - 1 - lo
            exchange_rows(off, rows, below, above, self.group)
            self._fetch_rows(st, ghost_rows(rows, self.rank, below, above, n))
            win = (C.c_int * 3)()
            st["out"][a0:a1].zero_()  # (the gather adds into its destination rows)
            _lib.call("bf_grid_regrid_gather_rows", dim, n, dv.ptr(off), dv.ptr(st["orig"]), dv.ptr(st["out"]), lohi, a0, a1,
                      win, dv.stream_ptr())
            self.regrid_window = tuple(win)
        return st["out"][a0:a1]

    def process(self, gather=True):
        from . import _device as dv
        r = self.runner
        G = r.GriddedMap
        dev = dv.device(self.device)
        with torch.cuda.device(dev):
            st = self.upload(dev)
            out = self.pairs_device(st)
            own = self.regrid_device(st, out) if self.baryonify else out[st["a0"]:st["a1"]]
            pairs = torch.tensor([self.local_pairs], dtype=torch.int64, device=dev)
            if self.world > 1:
                dist.all_reduce(pairs, group=self.group)
            self.npairs = int(pairs.item())
            if not gather:
                res = dv.to_host(own.contiguous(), True)
                if self.baryonify:  # Map2DRunner.py:611-613, over all slabs
                    sums = torch.stack([own.sum(), st["orig_slab"].sum()])
                    if self.world > 1:
                        dist.all_reduce(sums, group=self.group)
                    assert np.isclose(float(sums[0]), float(sums[1])), \
                        "ERROR in pixel regridding, sum(new_map) [%0.14e] != sum(oldmap) [%0.14e]" % (float(sums[0]), float(sums[1]))
                return res
            full = gather_slices(own.contiguous(), st["rows"], self.group)
            res = dv.to_host(full.reshape(-1), True).reshape(G.map.shape)
        if self.baryonify:  # Map2DRunner.py:611-613
            new_sum, old_sum = float(res.sum()), float(np.sum(G.map))
            assert np.isclose(new_sum, old_sum), \
                "ERROR in pixel regridding, sum(new_map) [%0.14e] != sum(oldmap) [%0.14e]" % (new_sum, old_sum)
        return res


def _runs(sorted_rows):
    """Consecutive runs [g0, g1) of a sorted list of row indices."""
    out, start, prev = [], None, None
    for g in sorted_rows:
        if start is None:
            start = prev = g
        elif g == prev + 1:
            prev = g
        else:
            out.append((start, prev + 1))
            start = prev = g
    if start is not None:
        out.append((start, prev + 1))
    return out


# ----------------------------------------------------------------------------------------------------------
# Delta P(k) on N GPUs: slab-decomposed transform (last row)
# ----------------------------------------------------------------------------------------------------------
def reduce_to_slab(full, rows, group=None):
    """Sum the ranks' full-size deposit grids [N, N, N] and keep this rank's x-slab ``rows[rank]``.  NGP counts are
    whole numbers, so the sum is exact in any order.  NCCL: one reduce-scatter when the slabs are equal, otherwise
    (and on gloo, which has no reduce-scatter) an all-reduce followed by the slice."""
    rank, world = _world(group)
    a, b = rows[rank]
    if world == 1:
        return full[a:b]
    equal = all(r[1] - r[0] == b - a for r in rows)
    if equal and dist.get_backend(group) == "nccl":
        out = torch.empty((b - a,) + tuple(full.shape[1:]), dtype=full.dtype, device=full.device)
        dist.reduce_scatter_tensor(out, full.contiguous(), group=group)
        return out
    dist.all_reduce(full, group=group)
    return full[a:b]


def slab_rfftn(slab, rows, group=None):
    """rfftn of a real [N, N, N] mesh held as x-slabs: rank g passes ``slab`` = planes ``rows[g]``.

    Local 2-D real transforms over (y, z) -> the only exchange of the path, an all-to-all TRANSPOSE (every rank
    sends rank h the y-range ``cols[h]`` of its planes and receives all x for its own y-range) -> 1-D transforms over
    x.  Returns ``(piece, (j0, j1))`` with ``piece[i, j - j0, l] = rfftn(mesh)[i, j, l]`` complex, all i, j in
    [j0, j1), all l = 0..N/2.  Device-agnostic (torch.fft is cuFFT on the GPU), point-to-point through
    ``batch_isend_irecv`` so that the same code runs on gloo in the CPU tests."""
    rank, world = _world(group)
    N = int(slab.shape[1])
    nh = N // 2 + 1
    cols = index_split(N, world)
    j0, j1 = cols[rank]
    A = torch.fft.rfft2(slab, dim=(1, 2))  # [nx_loc, N, nh]
    if world == 1:
        return torch.fft.fft(A, dim=0), (j0, j1)
    T = torch.empty((N, j1 - j0, nh), dtype=A.dtype, device=A.device)
    ops, keep = [], []
    for h in range(world):
        c0, c1 = cols[h]
        x0, x1 = rows[h]
        if h == rank:
            T[x0:x1] = A[:, j0:j1]
            continue
        if c1 > c0 and A.shape[0] > 0:
            payload = torch.view_as_real(A[:, c0:c1].contiguous())
            keep.append(payload)
            ops.append(dist.P2POp(dist.isend, payload, dist.get_global_rank(group, h) if group is not None else h,
                                  group=group))
        if x1 > x0 and j1 > j0:  # planes x0..x1 of T are one contiguous block: receive in place
            ops.append(dist.P2POp(dist.irecv, torch.view_as_real(T[x0:x1]),
                                  dist.get_global_rank(group, h) if group is not None else h, group=group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return torch.fft.fft(T, dim=0), (j0, j1)


class ShardedPowerSpectrum:
    """``pk.PowerSpectrumGrid`` on N ranks: every rank deposits ITS particles (any partition: e.g. the slice
    ``ShardedSnapshot.process(gather=False)`` returns), the mesh is reduced onto x-slabs, transformed with
    :func:`slab_rfftn`, k-binned piece by piece (``bf_pk_bin_part``) and the ``[3][Nk]`` bin sums are all-reduced, so
    every rank returns the whole spectrum.  Mode counts equal the single-GPU ones exactly; the power agrees to FFT
    rounding (the 2-D + 1-D factorisation adds in a different order than cuFFT's 3-D plan)."""

    def __init__(self, Ngrd, Lbox, Nk=180, group=None, device=None):
        from .pk import PowerSpectrumGrid
        self.local = PowerSpectrumGrid(Ngrd, Lbox, Nk, device=device)
        self.Ngrd, self.Lbox, self.Nk, self.group, self.device = int(Ngrd), float(Lbox), int(Nk), group, device
        self.rank, self.world = _world(group)
        self.rows = index_split(self.Ngrd, self.world)

    def density_slab(self, particles_local, factor=1):
        """This rank's x-slab of the NGP counts of ALL ranks' particles, [rows[rank], N, N] float64 on the device."""
        full = self.local.density(particles_local, factor)
        return reduce_to_slab(full, self.rows, self.group)

    def power_of_slab(self, slab, factor=1):
        from . import _device as dv
        dev = slab.device
        _, val, k0, dk = self.local._setup(factor)
        with torch.cuda.device(dev):
            piece, (j0, j1) = slab_rfftn(slab, self.rows, self.group)
            lib = _lib.load(require_cuda=True)
            ext = (C.c_int64 * 3)(self.Ngrd, j1 - j0, self.Ngrd // 2 + 1)
            off = (C.c_int64 * 3)(0, j0, 0)
            nb = lib.bf_pk_bin_part_ws_bytes(ext, self.Nk)
            ws = torch.empty(max(nb, 256), dtype=torch.uint8, device=dev)
            out = dv.empty((3, self.Nk), dev=dev)
            strides = (C.c_int64 * 3)(*piece.stride())
            _lib.call("bf_pk_bin_part", self.Ngrd, dv.ptr(piece) if piece.numel() else None, strides, ext, off, val, k0,
                      dk, self.Nk, dv.ptr(out), dv.ptr(ws), ws.numel(), dv.stream_ptr())
            if self.world > 1:
                dist.all_reduce(out, group=self.group)
            o = out.cpu().numpy()
        with np.errstate(divide="ignore", invalid="ignore"):
            return o[1] / o[2] * factor, o[0] / o[2], o[2]

    def power(self, particles_local, factor=1):
        """(k_cen * factor, P(k), modes per bin) of the union of the ranks' particle sets, on every rank."""
        return self.power_of_slab(self.density_slab(particles_local, factor), factor)

    def suppression(self, part_baryon_local, part_dmo_local, factors=(1, 8)):
        """{factor: {'k', 'S' = P_baryon / P_dmo}} as ``PowerSpectrumGrid.suppression``, on every rank."""
        res = {}
        pb, pd = self.local.upload(part_baryon_local), self.local.upload(part_dmo_local)
        for f in factors:
            k, PB, _ = self.power(pb, f)
            _, PD, _ = self.power(pd, f)
            res[f"Factor{f}"] = {"k": k, "S": PB / PD}
        return res
