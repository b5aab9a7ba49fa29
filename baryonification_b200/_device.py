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


def to_dev(a, dtype=np.float64, dev=None, pinned=False):
    """numpy -> device tensor (contiguous).  `pinned` stages through page-locked memory (async H2D)."""
    a = np.ascontiguousarray(a, dtype=dtype)
    t = torch.from_numpy(a)
    if pinned:
        t = t.pin_memory()
    return t.to(dev or device(), non_blocking=pinned)


class AsyncUpload:
    """Pageable numpy array -> HBM, off the critical path.

    A helper thread copies the caller's array into page-locked memory (torch's caching host allocator keeps the
    block for the next call; the CPU copy is multi-threaded) and enqueues the H2D copy on a side stream, while
    the main thread sets up halos and runs the binning / pair kernels.  ``result()`` joins and makes the current
    stream wait for the copy."""

    def __init__(self, a, dev, dtype=np.float64):
        import threading
        self.dev = dev
        self.side = torch.cuda.Stream(device=dev)
        self.done = torch.cuda.Event()
        self.out = None
        self.err = None
        a = np.ascontiguousarray(a, dtype=dtype)

        def work():
            try:
                with torch.cuda.device(dev):
                    src = torch.from_numpy(a) if a.flags.writeable else torch.from_numpy(a.copy())
                    pin = torch.empty(src.shape, dtype=src.dtype, pin_memory=True)
                    pin.copy_(src)
                    with torch.cuda.stream(self.side):
                        self.out = pin.to(dev, non_blocking=True)
                        self.done.record(self.side)
            except BaseException as e:  # re-raised in result()
                self.err = e

        self.th = threading.Thread(target=work, daemon=True)
        self.th.start()

    def result(self):
        self.th.join()
        if self.err is not None:
            raise self.err
        torch.cuda.current_stream(self.dev).wait_event(self.done)
        self.out.record_stream(torch.cuda.current_stream(self.dev))
        return self.out


class ChunkedUpload:
    """``AsyncUpload`` in consecutive pieces: ``edges`` cuts the (flat) array into ranges that are staged and copied in
    order, each with its own completion event, so that a consumer that works through the array front to back
    (the ring-band pipeline of ``ShellEngine.baryonify_to_host``) can start on the first range while the rest is still
    on its way.  ``wait(i)`` makes the current stream wait for range i; ``.out`` is the whole device array."""

    def __init__(self, a, dev, edges, dtype=np.float64):
        import threading
        self.dev = dev
        self.side = torch.cuda.Stream(device=dev)
        self.edges = [int(e) for e in edges]
        self.events = [torch.cuda.Event() for _ in range(len(self.edges) - 1)]
        self.ready = [threading.Event() for _ in self.events]
        self.err = None
        a = np.ascontiguousarray(a, dtype=dtype).reshape(-1)
        self.out = torch.empty(a.shape, dtype=torch.float64, device=dev)

        def work():
            try:
                with torch.cuda.device(dev):
                    src = torch.from_numpy(a) if a.flags.writeable else torch.from_numpy(a.copy())
                    pin = torch.empty(src.shape, dtype=src.dtype, pin_memory=True)
                    for i, (lo, hi) in enumerate(zip(self.edges[:-1], self.edges[1:])):
                        pin[lo:hi].copy_(src[lo:hi])
                        with torch.cuda.stream(self.side):
                            self.out[lo:hi].copy_(pin[lo:hi], non_blocking=True)
                            self.events[i].record(self.side)
                        self.ready[i].set()
                    self.pin = pin  # (alive until the copies are done; released with the object)
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
