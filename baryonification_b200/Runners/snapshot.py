"""Particle-snapshot runner on the GPU: ``BaryonifySnapshot``.

Drop-in for ``BaryonForge/Runners/SnapshotRunner.py`` (``DefaultRunnerSnapshot`` :21-155,
``BaryonifySnapshot`` :159-262).  ``KDTree_kwargs`` is accepted and ignored: the periodic ball query
is a cell-list gather on the device, there is no tree to build.
"""
from __future__ import annotations

import ctypes as C
import os
import time

import numpy as np
import torch

from .. import _device as dv
from .. import _lib, host
from .._ccl import default_mass_def
from ..tables import displacement_host_table

__all__ = ["DefaultRunnerSnapshot", "BaryonifySnapshot"]


class _RecordPipeline:
    """Particle records: pageable numpy -> pinned staging -> HBM -> kernel -> pinned result, chunk by chunk.

    PCIe is full duplex, so the upload of chunk i+1, the kernel on chunk i and the read-back of chunk i-1 overlap; a
    helper thread does the pageable -> pinned copies (started in the constructor, i.e. during halo set-up).  The
    result array is backed by page-locked memory from torch's caching host allocator."""

    CHUNK = int(os.environ.get("BF_SNAP_CHUNK", 1 << 20))  # particles per chunk (32 MiB of records)

    def __init__(self, rec, dev):
        import threading
        self.rec, self.dev = rec, dev
        n = rec.shape[0]
        self.bounds = [(i, min(i + self.CHUNK, n)) for i in range(0, n, self.CHUNK)]
        self.h2d, self.d2h = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        self.d_in = torch.empty((n, 4), dtype=torch.float64, device=dev)
        self.ready = [torch.cuda.Event() for _ in self.bounds]
        self.queued = [threading.Event() for _ in self.bounds]
        self.err = None
        self.consuming = threading.Event()
        self.aborted = False
        self.th = threading.Thread(target=self._upload, daemon=True)
        self.th.start()

    def _upload(self):
        try:
            import warnings
            dv.thread_budget()  # (a helper thread does not inherit the main thread's torch.set_num_threads)
            with warnings.catch_warnings():  # a read-only input is only ever read here
                warnings.simplefilter("ignore")
                src = torch.from_numpy(self.rec)
            with torch.cuda.device(self.dev):
                if dv.pin_in_place(self.rec):  # page-locked where it lies (a repeated input): no staging copy
                    # Nothing to prepare on the host, so the records only start to travel when run() consumes them: the
                    # halo set-up's own (small, partly synchronous) uploads share the GPU's one host-to-device copy
                    # engine and stood behind the records for the whole 78 ms of their transfer (measured: set-up 84
                    # instead of 6 ms, 160 instead of 121 ms end to end).
                    self.consuming.wait()
                    if self.aborted:
                        return
                    for i, (a, b) in enumerate(self.bounds):
                        with torch.cuda.stream(self.h2d):
                            self.d_in[a:b].copy_(src[a:b], non_blocking=True)
                            self.ready[i].record(self.h2d)
                        self.queued[i].set()
                    return
                stage = [torch.empty((min(self.CHUNK, src.shape[0]), 4), dtype=torch.float64, pin_memory=True)
                         for _ in range(3)]
                free = [None] * 3
                for i, (a, b) in enumerate(self.bounds):
                    k = i % 3
                    if free[k] is not None:
                        free[k].synchronize()  # the H2D copy that last used this staging buffer is done
                    stage[k][: b - a].copy_(src[a:b])  # (torch's OpenMP copy: a Python thread pool per 32 MB chunk
                    # fights the main thread for the GIL -- measured 181 vs 138 ms end to end)
                    with torch.cuda.stream(self.h2d):
                        self.d_in[a:b].copy_(stage[k][: b - a], non_blocking=True)
                        self.ready[i].record(self.h2d)
                        free[k] = self.ready[i]
                    self.queued[i].set()
        except BaseException as e:  # re-raised in run()
            self.err = e
            for q in self.queued:
                q.set()

    def abort(self):
        """The caller gives up before ``run()`` (its set-up raised): let the upload thread finish and drop the buffers."""
        self.aborted = True
        self.consuming.set()
        self.th.join()
        self.d_in = None

    def run(self, kernel, out=None):
        """kernel(rec_chunk, out_chunk, stream) per chunk on the current stream; returns the host result array.
        ``out``: a page-locked [n, 4] float64 host tensor to receive it (a rank's piece of a result array the ranks of a
        box share); default: a fresh one from torch's caching host allocator."""
        n = self.rec.shape[0]
        cur = torch.cuda.current_stream(self.dev)
        d_out = torch.empty((n, 4), dtype=torch.float64, device=self.dev)
        if out is None:
            out = torch.empty((n, 4), dtype=torch.float64, pin_memory=True)
        assert tuple(out.shape) == (n, 4) and out.dtype == torch.float64 and out.is_contiguous()
        done = torch.cuda.Event()
        self.consuming.set()
        for i, (a, b) in enumerate(self.bounds):
            self.queued[i].wait()
            if self.err is not None:
                raise self.err
            cur.wait_event(self.ready[i])
            kernel(self.d_in[a:b], d_out[a:b], cur)
            done = torch.cuda.Event()
            done.record(cur)
            self.d2h.wait_event(done)
            with torch.cuda.stream(self.d2h):
                out[a:b].copy_(d_out[a:b], non_blocking=True)
        self.th.join()
        self.d2h.synchronize()
        cur.synchronize()
        return out.numpy()


class DefaultRunnerSnapshot:
    def __init__(self, HaloNDCatalog, ParticleSnapshot, epsilon_max, model, mass_def=None, verbose=True,
                 KDTree_kwargs={}, *, device=None):
        self.HaloNDCatalog = HaloNDCatalog
        self.ParticleSnapshot = ParticleSnapshot
        self.epsilon_max = epsilon_max
        self.cosmo = HaloNDCatalog.cosmology
        self.model = model
        self.mass_def = mass_def if mass_def is not None else default_mass_def()
        self.verbose = verbose
        self.device = device
        self.npairs = None
        self.timings = {}


    def _progress(self, what):
        """Per-stage progress lines under the reference's ``verbose`` flag (its tqdm bar over halos,
        SnapshotRunner.py:205, has no per-halo counterpart here)."""
        if self.verbose:
            import sys
            print("[%s] %s" % (type(self).__name__, what), file=sys.stderr, flush=True)


class BaryonifySnapshot(DefaultRunnerSnapshot):
    """``process()`` returns a displaced copy of ``ParticleSnapshot.cat`` (SnapshotRunner.py:250-262)."""

    @staticmethod
    def _is_reference_layout(cat):
        """(M, x, y, z) native float64 records of 32 bytes (utils/io.py:581-589): the kernel reads them as they are."""
        d = cat.dtype
        if d.names is None or d.itemsize != 32 or not cat.flags.c_contiguous:
            return False
        want = {"M": 0, "x": 8, "y": 16, "z": 24}
        return all(k in d.fields and d.fields[k][1] == o and d.fields[k][0] == np.dtype("=f8") for k, o in want.items())

    def _upload(self, dev):
        """Host part of the set-up: per-halo scalars (vectorised numpy) and their H2D copies, table upload."""
        S = self.ParticleSnapshot
        if len(vars(self.model).get("p_keys", [])) > 0:
            raise NotImplementedError("property-keyed tables are not yet supported by the B200 runners")
        lib = _lib.load(require_cuda=True)
        table = dv.DeviceTable(displacement_host_table(self.model), dev=dev)
        s = host.snapshot_scalars(self.HaloNDCatalog, S, self.mass_def, self.epsilon_max, self.model)
        dim, nh = s["dim"], s["Rq"].size
        t = {k: dv.to_dev(s[k], dev=dev) for k in ("pos", "Rq", "cut", "lnr_shift", "z_in", "M_in")}
        L = float(S.L)
        ng = lib.bf_snap_grid_cells(dim, L, float(s["Rq"].mean()) if nh else 0.0)
        if os.environ.get("BF_SNAP_NG"):  # tuning experiments only
            ng = int(os.environ["BF_SNAP_NG"])
        return dict(table=table, s=s, t=t, ng=ng, dim=dim, L=L, nh=nh, ws=None, max_entries=0)

    def _bin(self, st, dev):
        """Device part of the set-up: (z, M) collapse and halo binning into the periodic grid.  The workspace of an
        earlier call on the same ``st`` is reused (the counting pass only runs when it does not fit)."""
        lib = _lib.load(require_cuda=True)
        t, dim, nh, L, ng = st["t"], st["dim"], st["nh"], st["L"], st["ng"]
        st["rows"] = rows = st["table"].collapse(t["z_in"], t["M_in"])
        st["desc"] = desc = _lib.SnapHalos(t["pos"].data_ptr(), t["Rq"].data_ptr(), t["cut"].data_ptr(),
                                           t["lnr_shift"].data_ptr(), rows.data_ptr(), nh)
        for _ in range(24):  # counting pass first (exact size), then the real one; a coarser grid when the list is too long
            ng = st["ng"]
            need = max(lib.bf_snap_ws_bytes(dim, nh, L, ng, st["max_entries"]), 256)
            if st["ws"] is None or st["ws"].numel() < need:
                st["ws"] = None
                st["ws"] = torch.empty(need, dtype=torch.uint8, device=dev)
            n = C.c_int64()
            rc = lib.bf_snap_bin_halos(dim, L, ng, C.byref(desc), st["max_entries"], dv.ptr(st["ws"]), st["ws"].numel(),
                                       C.byref(n), dv.stream_ptr())
            if rc == -2 and n.value > st["max_entries"]:
                st["max_entries"] = n.value
                continue
            if rc == -4 and ng > 1:  # BF_E_CAPACITY: >= 2^25 (cell, halo) entries; the list shrinks ~ ng^dim
                st["ng"], st["max_entries"] = max(1, ng // 2), 0
                continue
            _lib.check(rc, "bf_snap_bin_halos")
            break
        else:
            raise _lib.BFError("bf_snap_bin_halos: workspace retry failed")
        self.nentries = n.value
        return st

    def _setup(self, dev):
        """Table, per-halo arrays, (z, M) collapse and halo binning into the periodic grid."""
        return self._bin(self._upload(dev), dev)

    def pair_list(self):
        """The (particle, halo) pairs the run evaluates -- the union of the reference's per-halo
        ``tree.query_ball_point`` answers (SnapshotRunner.py:219 / :234) -- as a sorted uint64 array of
        ``particle << 32 | halo``.  A diagnostic for the parity tests; ``process()`` does not build it."""
        dev = dv.device(self.device)
        cat = self.ParticleSnapshot.cat
        with torch.cuda.device(dev):
            st = self._setup(dev)
            dim, s = st["dim"], st["s"]
            names = ("x", "y", "z")[:dim]
            xin = [dv.to_dev(cat[k], dev=dev) for k in names]
            xout = [torch.empty_like(v) for v in xin]
            count = torch.zeros(1, dtype=torch.int64, device=dev)
            cap = 0
            for _ in range(2):  # first pass counts, second pass stores
                pairs = torch.empty(max(cap, 1), dtype=torch.int64, device=dev)
                _lib.call("bf_snap_pair_list", dim, xin[0].numel(), dv.ptr(xin[0]), dv.ptr(xin[1]),
                          dv.ptr(xin[2] if dim == 3 else None), st["L"], float(s["a"]), C.byref(st["desc"]),
                          C.byref(st["table"].desc), st["ng"], dv.ptr(st["ws"]), dv.ptr(xout[0]), dv.ptr(xout[1]),
                          dv.ptr(xout[2] if dim == 3 else None), dv.ptr(pairs), cap, dv.ptr(count), dv.stream_ptr())
                n = int(count.item())
                if n <= cap:
                    break
                cap = n
            self.npairs = n
            return np.sort(pairs[:n].cpu().numpy().view(np.uint64))

    def process(self, _out=None):
        """``_out`` (internal, ``sharding.ShardedSnapshot``): page-locked [n, 4] float64 host tensor that receives the
        displaced (M, x, y, z) records instead of a fresh array; only honoured for the reference's record layout."""
        dev = dv.device(self.device)
        S = self.ParticleSnapshot
        with torch.cuda.device(dev):
            t0 = time.perf_counter()
            cat = S.cat
            fast = self._is_reference_layout(cat)
            pipe = _RecordPipeline(cat.view(np.float64).reshape(-1, 4), dev) if fast else None  # travels during set-up
            self._progress("host set-up of %d halos, halo binning into the periodic grid" % len(self.HaloNDCatalog.cat))
            try:
                st = self._setup(dev)
            except BaseException:
                if pipe is not None:  # release the upload thread (it may be waiting for run()) and its device buffer
                    pipe.abort()
                raise
            self._progress("ball query + displacement of %d particles (upload / kernel / read-back in chunks)" % cat.size)
            dim, s = st["dim"], st["s"]
            cnt = dv.pair_counter(dev)
            self.timings["setup"] = time.perf_counter() - t0
            t1 = time.perf_counter()
            if fast:
                def displace(rec, rec_out, stream):
                    _lib.call("bf_snap_displace_records", dim, rec.shape[0], dv.ptr(rec), st["L"], float(s["a"]),
                              C.byref(st["desc"]), C.byref(st["table"].desc), st["ng"], dv.ptr(st["ws"]),
                              dv.ptr(rec_out), dv.ptr(cnt), C.c_void_p(stream.cuda_stream))
                new_cat = pipe.run(displace, out=_out).reshape(-1).view(cat.dtype).reshape(cat.shape)
            else:  # any other record layout: field by field
                names = ("x", "y", "z")[:dim]
                xin = [dv.to_dev(cat[k], dev=dev, pinned=True) for k in names]
                xout = [torch.empty_like(v) for v in xin]
                _lib.call("bf_snap_displace", dim, xin[0].numel(), dv.ptr(xin[0]), dv.ptr(xin[1]),
                          dv.ptr(xin[2] if dim == 3 else None), st["L"], float(s["a"]), C.byref(st["desc"]),
                          C.byref(st["table"].desc), st["ng"], dv.ptr(st["ws"]), dv.ptr(xout[0]), dv.ptr(xout[1]),
                          dv.ptr(xout[2] if dim == 3 else None), dv.ptr(cnt), dv.stream_ptr())
                new_cat = cat.copy()  # SnapshotRunner.py:250
                for k, v in zip(names, xout):
                    new_cat[k] = v.cpu().numpy()
            self.npairs = int(cnt.item())
            self.timings["displace_d2h"] = time.perf_counter() - t1
            self.timings["process"] = time.perf_counter() - t0
        return new_cat
