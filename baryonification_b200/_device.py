"""Device plumbing: torch owns allocations / streams; kernels see raw pointers through the C ABI."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .tables import HostTable


def device(index=None):
    _lib.load(require_cuda=True)
    if index is None:
        index = torch.cuda.current_device()
    return torch.device("cuda", index)


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return C.c_void_p(0) if t is None else C.c_void_p(t.data_ptr())


def thread_budget():
    """Call first in every helper thread that copies host memory with torch.  torchrun starts the ranks with
    OMP_NUM_THREADS=1, and a thread created later takes its OpenMP team size from that environment default, not from the
    main thread's ``torch.set_num_threads`` -- its pageable -> page-locked staging copies (GBs per call) would run on one
    core.  Gives the thread this rank's share of the host cores (never fewer than it has)."""
    from .host import host_cores
    n = host_cores()
    if torch.get_num_threads() < n:
        torch.set_num_threads(n)
    return n


_copy_pool = None


def host_copy(dst, src):
    """``dst.copy_(src)`` for contiguous host tensors of one dtype, as ``numpy.copyto`` of equal pieces in a thread pool
    of this rank's share of the host cores (numpy releases the GIL in its copy loops).  Measured on the B200 boxes
    (scripts/host_copy_probe.py, 800 MB pageable -> page-locked): torch ``copy_`` 7.7 GB/s on one thread -- what a helper
    thread gets under torchrun's OMP_NUM_THREADS=1 -- and 46 GB/s on 8; numpy in a pool 15.7 GB/s on one thread, 75 GB/s
    on 8; the PCIe link moves 55 GB/s, so only the pooled copy keeps the staging off the critical path."""
    global _copy_pool
    n = dst.numel()
    if (n < (1 << 20) or dst.dtype != src.dtype or not dst.is_contiguous() or not src.is_contiguous()
            or dst.device.type != "cpu" or src.device.type != "cpu"):
        dst.copy_(src)
        return dst
    from concurrent.futures import ThreadPoolExecutor
    from .host import host_cores
    nt = max(1, min(16, host_cores()))
    if _copy_pool is None or _copy_pool._max_workers != nt:
        _copy_pool = ThreadPoolExecutor(max_workers=nt, thread_name_prefix="bf-copy")
    import warnings
    with warnings.catch_warnings():  # (a read-only source is only ever read)
        warnings.simplefilter("ignore")
        d, s = dst.view(-1).numpy(), src.detach().view(-1).numpy()
    edges = np.linspace(0, n, nt + 1).astype(np.int64)
    list(_copy_pool.map(lambda i: np.copyto(d[edges[i]:edges[i + 1]], s[edges[i]:edges[i + 1]]), range(nt)))
    return dst


# ---- page-locking a caller's array in place -------------------------------------------------------------------------
# Staging a large input through page-locked memory costs a host memcpy that all ranks of a box share (1.6 GB of map at
# ~76 GB/s = 21 ms however many GPUs work on it) -- hidden behind a 54 ms pair kernel on one GPU, exposed behind a 7 ms
# one on eight.  Baryonification runs are parameter sweeps: the same map / snapshot is processed again and again with
# other models.  So an input of >= 64 MiB is staged the FIRST time it is seen and page-locked where it lies
# (cudaHostRegister) the SECOND time; from then on its rows go to the GPU straight from the caller's memory at PCIe
# speed.  One-shot callers never pay the registration (~0.1 ms per MB), and the registration ends with the array
# (weakref finaliser on the array that owns the memory).  BF_PIN_INPUTS=0 turns it off, =1 registers at first sight.
_PIN_MIN_BYTES = 64 << 20
_pin_seen, _pin_done = {}, {}


def _root_owner(a):
    while isinstance(getattr(a, "base", None), np.ndarray):
        a = a.base
    return a


def _unpin(key):
    ptr = _pin_done.pop(key, None)
    if ptr is not None:
        try:
            _lib.load().bf_host_unpin(C.c_void_p(ptr))
        except Exception:
            pass


def pin_in_place(a):
    """True when the memory of the C-contiguous numpy array ``a`` is page-locked (now or already): the caller may issue
    asynchronous copies straight from it."""
    import os
    import weakref
    mode = os.environ.get("BF_PIN_INPUTS", "")
    if mode == "0" or not torch.cuda.is_available() or not a.flags.c_contiguous or a.nbytes < _PIN_MIN_BYTES:
        return False
    root = _root_owner(a)
    lo = a.__array_interface__["data"][0]
    for (p0, n0) in list(_pin_done):  # (a slice of a registered range)
        if p0 <= lo and lo + a.nbytes <= p0 + n0:
            return True
    key = (lo, a.nbytes)  # (the range that is used -- a rank's slice of the map -- not the whole owning array)
    state = _pin_seen.get(key)
    if state == "failed":
        return False
    if state == "pinned":  # page-locked by its owner (a Runner's own result array fed back in)
        return True
    if state is None and mode != "1":
        _pin_seen[key] = "seen"
        try:  # forget the sighting when the array goes away (its address may be reused)
            weakref.finalize(root, _pin_seen.pop, key, None)
        except TypeError:
            _pin_seen[key] = "failed"
        return False
    try:
        fin = weakref.finalize(root, _unpin, key)
    except TypeError:
        _pin_seen[key] = "failed"
        return False
    rc = int(_lib.load(require_cuda=True).bf_host_pin(C.c_void_p(lo), a.nbytes))
    if rc != 0:
        fin.detach()
        _pin_seen[key] = "pinned" if rc == 1 else "failed"
        return rc == 1
    _pin_done[key] = lo
    return True


def pinned_already(a):
    """True when ``a`` lies in memory that ``pin_in_place`` has page-locked before (no registration is attempted): an
    upload of it needs no host-side preparation, so it need not start early."""
    if not isinstance(a, np.ndarray) or not a.flags.c_contiguous:
        return False
    lo = a.__array_interface__["data"][0]
    if _pin_seen.get((lo, a.nbytes)) == "pinned":
        return True
    return any(p0 <= lo and lo + a.nbytes <= p0 + n0 for (p0, n0) in list(_pin_done))


def _as_tensor(a):
    import warnings
    with warnings.catch_warnings():  # (a read-only input is only ever read)
        warnings.simplefilter("ignore")
        return torch.from_numpy(a)


def to_dev(a, dtype=np.float64, dev=None, pinned=False):
    """numpy -> device tensor (contiguous).  `pinned` stages through page-locked memory (async H2D) unless the array is
    page-locked in place (``pin_in_place``)."""
    a = np.ascontiguousarray(a, dtype=dtype)
    t = _as_tensor(a)
    if pinned and not pin_in_place(a):
        t = host_copy(torch.empty(t.shape, dtype=t.dtype, pin_memory=True), t)
    return t.to(dev or device(), non_blocking=pinned)


_PIECE = 8 << 20  # elements (64 MiB of float64) per H2D request of a large upload


def _trickle(dev, side, src, out, pinned, edges, on_range):
    """Copy the flat host tensor ``src`` to the device tensor ``out`` range by range (``edges``) on stream ``side``,
    calling ``on_range(i)`` (with ``side`` current) when range i has been enqueued.  Every range goes as requests of at
    most 64 MiB, and a request is only enqueued when the one before it has finished: the GPU has ONE host-to-device copy
    engine, and the per-halo scalars, tables and counters that the other streams upload meanwhile would otherwise queue
    behind the whole map (measured: 30 ms added to the binning stage when 1.6 GB was enqueued at once).  With ``pinned``
    false every request is staged through page-locked memory first (``host_copy``)."""
    half = max(1, min(_PIECE, max((hi - lo for lo, hi in zip(edges[:-1], edges[1:])), default=1)))  # largest request
    stage = None if pinned else torch.empty(2 * half, dtype=src.dtype, pin_memory=True)
    prev, k = [None, None], 0
    for i, (lo, hi) in enumerate(zip(edges[:-1], edges[1:])):
        for p0 in range(lo, hi, _PIECE):
            p1 = min(p0 + _PIECE, hi)
            slot = k & 1
            if prev[slot] is not None:
                prev[slot].synchronize()  # (also: the staging half this request reuses is free again)
            if pinned:
                piece = src[p0:p1]
                if prev[slot ^ 1] is not None:
                    prev[slot ^ 1].synchronize()
            else:
                piece = host_copy(stage[slot * half: slot * half + (p1 - p0)], src[p0:p1])
            with torch.cuda.stream(side):
                out[p0:p1].copy_(piece, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(side)
            prev[slot] = ev
            k += 1
        with torch.cuda.stream(side):
            on_range(i)
    return stage


class ChunkedUpload:
    """Pageable (or page-locked-in-place, ``pin_in_place``) numpy array -> HBM, off the critical path, in consecutive
    ranges: ``edges`` cuts the (flat) array into ranges that travel in order, each with its own completion event, so that
    a consumer that works through the array front to back (the ring-band pipeline of ``ShellEngine.baryonify_to_host``)
    can start on the first range while the rest is still on its way.  A helper thread feeds the copy engine
    (``_trickle``) while the main thread sets up halos and runs the binning / pair kernels.  ``wait(i)`` makes the
    current stream wait for range i; ``result()`` for all of them; ``.out`` is the whole device array."""

    def __init__(self, a, dev, edges, dtype=np.float64):
        import threading
        self.dev = dev
        self.side = torch.cuda.Stream(device=dev)
        self.edges = [int(e) for e in edges]
        self.events = [torch.cuda.Event() for _ in range(len(self.edges) - 1)]
        self.ready = [threading.Event() for _ in self.events]
        self.err = None
        self.keep = a = np.ascontiguousarray(a, dtype=dtype).reshape(-1)
        self.out = torch.empty(a.shape, dtype=torch.float64, device=dev)
        pinned = pin_in_place(a)

        def on_range(i):
            self.events[i].record(self.side)
            self.ready[i].set()

        def work():
            try:
                thread_budget()
                with torch.cuda.device(dev):
                    self.stage = _trickle(dev, self.side, _as_tensor(a), self.out, pinned, self.edges, on_range)
            except BaseException as e:  # re-raised in wait()
                self.err = e
                for r in self.ready:
                    r.set()

        self.th = threading.Thread(target=work, daemon=True)
        self.th.start()

    def wait(self, i):
        self.ready[i].wait()
        if self.err is not None:
            raise self.err
        torch.cuda.current_stream(self.dev).wait_event(self.events[i])

    def result(self):
        for i in range(len(self.events)):
            self.wait(i)
        self.th.join()
        self.out.record_stream(torch.cuda.current_stream(self.dev))
        return self.out


class AsyncUpload(ChunkedUpload):
    """``ChunkedUpload`` of the whole array as one range (any shape)."""

    def __init__(self, a, dev, dtype=np.float64):
        a = np.ascontiguousarray(a, dtype=dtype)
        self.shape = a.shape
        super().__init__(a, dev, [0, a.size], dtype)

    def result(self):
        return super().result().view(self.shape)


def to_host(t, pinned=True):
    """Device tensor -> numpy array owned by the caller.  With ``pinned`` the array is backed by page-locked
    memory from torch's caching host allocator (returned to the cache when the array is garbage-collected): the
    D2H copy then runs at PCIe speed instead of the pageable-staging speed."""
    if not pinned:
        return t.cpu().numpy()
    out = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
    out.copy_(t, non_blocking=True)
    torch.cuda.current_stream(t.device).synchronize()
    return out.numpy()


def zeros(shape, dtype=torch.float64, dev=None):
    return torch.zeros(shape, dtype=dtype, device=dev or device())


def empty(shape, dtype=torch.float64, dev=None):
    return torch.empty(shape, dtype=dtype, device=dev or device())


class DeviceTable:
    """A HostTable resident in HBM + the bf_table descriptor pointing at it."""

    def __init__(self, host: HostTable, dev=None):
        self.host = host
        self.gz, self.gM, self.gr = (to_dev(x, dev=dev) for x in (host.gz, host.gM, host.gr))
        self.values = to_dev(host.values, dev=dev)
        self.desc = _lib.Table(self.gz.data_ptr(), self.gM.data_ptr(), self.gr.data_ptr(), self.values.data_ptr(),
                               host.gz.size, host.gM.size, host.gr.size, float(host.gr[0]), float(host.gr[-1]))
        self.nr = host.gr.size
        self.nd = None
        if host.extra:  # extra property axes (p_keys): N-d collapse descriptor
            self.gx = [to_dev(g, dev=dev) for g in host.extra]
            grids = [self.gz, self.gM] + self.gx
            nd = _lib.TableND()
            nd.naxes = len(grids)
            for a, g in enumerate(grids):
                nd.grid[a] = g.data_ptr()
                nd.n[a] = g.numel()
            nd.values = self.values.data_ptr()
            nd.nr = self.nr
            self.nd = nd

    def collapse(self, z_in, M_in, extras=None):
        """rows[j, :] for every halo (bf_table_collapse_zM; bf_table_collapse_nd when the table has property axes,
        ``extras`` = [nh, n_keys] per-halo property values)."""
        nh = z_in.numel()
        rows = empty((nh, self.nr), dev=z_in.device)
        if self.nd is None:
            _lib.call("bf_table_collapse_zM", C.byref(self.desc), nh, ptr(z_in), ptr(M_in), ptr(rows), stream_ptr())
        else:
            if extras is None or extras.shape != (nh, len(self.gx)):
                raise _lib.BFError("this table has property axes: per-halo property values [nh, n_keys] are required")
            coords = torch.cat([z_in.reshape(-1, 1), M_in.reshape(-1, 1), extras], dim=1).contiguous()
            _lib.call("bf_table_collapse_nd", C.byref(self.nd), nh, ptr(coords), ptr(rows), stream_ptr())
        return rows

    def readout(self, row, r, lnr_shift=0.0, cut=0.0, mode=0):
        out = torch.empty_like(r)
        _lib.call("bf_table_readout", C.byref(self.desc), ptr(row), r.numel(), ptr(r), float(lnr_shift), float(cut),
                  int(mode), ptr(out), stream_ptr())
        return out


def pair_counter(dev=None):
    return torch.zeros(1, dtype=torch.int64, device=dev or device())
