"""Device-resident execution plans ("engines") behind the Runner classes and the benchmark.

An engine owns the HBM-resident inputs of one runner call (per-halo arrays, table, collapsed rows, binning
workspace) and exposes the kernel stages as methods, so that ``Runner.process()`` (host buffers in, host
buffers out) and ``bench.py`` (inputs already resident) drive exactly the same code.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _device as dv
from . import _lib
from .tables import HostTable

PAINT, BARYONIFY = 0, 1


def ring_first_pixel(nside, ring):
    """First RING-scheme pixel of ring ``ring`` (1 .. 4 nside - 1); 12 nside^2 for ring 4 nside."""
    ns, ring = int(nside), int(ring)
    if ring < ns:
        return 2 * ring * (ring - 1)
    if ring < 3 * ns:
        return 2 * ns * (ns - 1) + (ring - ns) * 4 * ns
    k = 4 * ns - ring
    return 12 * ns * ns - 2 * k * (k + 1)


def band_groups(nside, ngroups):
    """``ngroups`` runs of ring bands with (nearly) equal pixel counts and their pixel edges: the units of the pipelined
    runs (pair stage, re-binning and read-back proceed group by group)."""
    from .sharding import band_pixel_range, band_split
    nb = int(_lib.load().bf_shell_num_bands(int(nside)))
    groups = [g for g in band_split(nside, max(1, min(int(ngroups), nb))) if g[1] > g[0]]
    edges = [band_pixel_range(nside, groups[0][0], groups[0][1])[0]] + [band_pixel_range(nside, lo, hi)[1] for lo, hi in groups]
    return groups, edges


class ShellEngine:
    def __init__(self, nside, mode, table: HostTable, scalars: dict, dev, variant="gather", bands=None):
        self.nside, self.mode, self.dev, self.variant = int(nside), int(mode), dev, variant
        self.lib = _lib.load(require_cuda=True)
        self.npix = 12 * self.nside ** 2
        self.ncomp = 1 if mode == PAINT else 3
        nb = self.lib.bf_shell_num_bands(self.nside)
        self.bands = (0, nb) if bands is None else (int(bands[0]), int(bands[1]))
        f, l = C.c_int64(), C.c_int64()
        _lib.check(self.lib.bf_shell_band_pixel_range(self.nside, self.bands[0], self.bands[1], C.byref(f), C.byref(l)))
        self.pix_first, self.pix_last = f.value, l.value
        self.host_table = table
        self.s = scalars
        self.nh = int(np.asarray(scalars["a"]).size) if scalars is not None else 0  # (upload_device sets it)
        self.table = None
        self.ws = None
        self.max_entries = 0  # size hint of the (cell, halo) key list: 0 = find out with a counting pass
        self.nentries = None
        self.npairs = None

    # ---- stage 0: H2D + (z, M) collapse -----------------------------------------------------
    def upload(self, pinned=False):
        dev, s = self.dev, self.s
        self.table = dv.DeviceTable(self.host_table, dev=dev)
        keys = ("vec", "lon", "lat", "radius", "D", "a", "cut", "lnr_shift", "z_in", "M_in")
        pins = s.get("_pinned") or {}  # page-locked outputs of the parallel host set-up: asynchronous copies
        self.t = {k: pins[k].to(dev, non_blocking=True) if k in pins else dv.to_dev(s[k], dev=dev, pinned=pinned)
                  for k in keys}
        self.t["extras"] = dv.to_dev(s["extras"], dev=dev) if s.get("extras") is not None else None  # p_keys columns
        self.rows = self.table.collapse(self.t["z_in"], self.t["M_in"], self.t["extras"])
        self.order = torch.argsort(self.t["radius"], descending=True) if self.variant == "atomic" else None
        t = self.t
        self.desc = _lib.ShellHalos(t["vec"].data_ptr(), t["lon"].data_ptr(), t["lat"].data_ptr(),
                                    t["radius"].data_ptr(), t["D"].data_ptr(), t["a"].data_ptr(),
                                    t["cut"].data_ptr(), t["lnr_shift"].data_ptr(), self.rows.data_ptr(),
                                    self.order.data_ptr() if self.order is not None else None, self.nh)
        self.cnt = dv.pair_counter(dev)
        return self

    # ---- stage 0 + 1 for a catalogue that arrives in runs of consecutive halos (host.chunked, runs > 1) -----------------
    RUN_KEYS = ("vec", "lon", "lat", "radius", "D", "a", "cut", "lnr_shift", "z_in", "M_in")

    def begin_runs(self, nh):
        """Device arrays for ``nh`` halos, table, binning workspace for ``self.max_entries`` keys (a size hint from an
        earlier call: without one the key list cannot be sized before the halos are known, use upload() + bin())."""
        assert self.variant == "gather" and self.max_entries > 0
        dev = self.dev
        self.nh = int(nh)
        self.table = dv.DeviceTable(self.host_table, dev=dev)
        self.t = {k: dv.empty((self.nh, 3) if k == "vec" else (self.nh,), dev=dev) for k in self.RUN_KEYS}
        self.t["extras"] = None
        self.rows = dv.empty((self.nh, self.table.nr), dev=dev)
        self.order = None
        t = self.t
        self.desc = _lib.ShellHalos(t["vec"].data_ptr(), t["lon"].data_ptr(), t["lat"].data_ptr(),
                                    t["radius"].data_ptr(), t["D"].data_ptr(), t["a"].data_ptr(),
                                    t["cut"].data_ptr(), t["lnr_shift"].data_ptr(), self.rows.data_ptr(), None, self.nh)
        self.cnt = dv.pair_counter(dev)
        need = self.lib.bf_shell_bin_ws_bytes(self.nside, self.nh, self.max_entries)
        if self.ws is None or self.ws.numel() < need:
            self.ws = None
            self.ws = torch.empty(need, dtype=torch.uint8, device=dev)
        _lib.call("bf_shell_bin_begin", self.nside, self.mode, C.byref(self.desc), self.bands[0], self.bands[1],
                  self.max_entries, dv.ptr(self.ws), self.ws.numel(), dv.stream_ptr())
        return self

    def add_run(self, j0, j1, pins):
        """Halos j0 <= j < j1 are ready in the page-locked tensors ``pins``: upload them, collapse their table rows, emit
        their (cell, halo) keys -- all asynchronous, the host goes on with the next run."""
        if j1 <= j0:
            return
        with torch.cuda.device(self.dev):
            for k in self.RUN_KEYS:
                self.t[k][j0:j1].copy_(pins[k][j0:j1], non_blocking=True)
            _lib.call("bf_table_collapse_zM", C.byref(self.table.desc), j1 - j0, dv.ptr(self.t["z_in"][j0:j1]),
                      dv.ptr(self.t["M_in"][j0:j1]), dv.ptr(self.rows[j0:j1]), dv.stream_ptr())
            _lib.call("bf_shell_bin_add", self.nside, self.mode, C.byref(self.desc), j0, j1, self.bands[0], self.bands[1],
                      self.max_entries, dv.ptr(self.ws), self.ws.numel(), dv.stream_ptr())

    def end_runs(self):
        """Sort the keys and mark the cell boundaries; a key list that outgrew the hint is redone with the exact size."""
        n = C.c_int64()
        rc = self.lib.bf_shell_bin_end(self.nside, self.mode, C.byref(self.desc), self.bands[0], self.bands[1],
                                       self.max_entries, dv.ptr(self.ws), self.ws.numel(), C.byref(n), dv.stream_ptr())
        if rc == -2 and n.value > self.max_entries:
            self.max_entries = n.value
            return self.bin()
        _lib.check(rc, "bf_shell_bin_end")
        self.nentries = n.value
        return self

    def upload_device(self, t, order=None):
        """``upload()`` for per-halo arrays that are already on the device (``sharding.ShardedShell``: prepared 1/N per
        rank, all-gathered over NVLink).  ``t``: dict of float64 tensors with the keys of ``upload()`` (+ "extras")."""
        dev = self.dev
        self.table = dv.DeviceTable(self.host_table, dev=dev)
        self.t = dict(t)
        self.nh = int(self.t["a"].numel())
        self.rows = self.table.collapse(self.t["z_in"], self.t["M_in"], self.t.get("extras"))
        self.order = torch.argsort(self.t["radius"], descending=True) if self.variant == "atomic" else None
        t = self.t
        self.desc = _lib.ShellHalos(t["vec"].data_ptr(), t["lon"].data_ptr(), t["lat"].data_ptr(),
                                    t["radius"].data_ptr(), t["D"].data_ptr(), t["a"].data_ptr(),
                                    t["cut"].data_ptr(), t["lnr_shift"].data_ptr(), self.rows.data_ptr(),
                                    self.order.data_ptr() if self.order is not None else None, self.nh)
        self.cnt = dv.pair_counter(dev)
        return self

    def recollapse(self):
        self.rows = self.table.collapse(self.t["z_in"], self.t["M_in"], self.t["extras"])
        self.desc.rows = self.rows.data_ptr()

    # ---- stage 1: halo binning into sky cells (gather variant) ----------------------------------
    def bin(self):
        if self.variant != "gather":
            return self
        # first call: a counting pass with a key-less workspace returns the exact number of (cell, halo) entries
        # (BF_E_WORKSPACE + *nentries_host), the second call is sized for it; later calls reuse the workspace
        for _ in range(2):
            need = self.lib.bf_shell_bin_ws_bytes(self.nside, self.nh, self.max_entries)
            if self.ws is None or self.ws.numel() < need:
                self.ws = None
                self.ws = torch.empty(need, dtype=torch.uint8, device=self.dev)
            n = C.c_int64()
            rc = self.lib.bf_shell_bin_halos(self.nside, self.mode, C.byref(self.desc), self.bands[0], self.bands[1],
                                             self.max_entries, dv.ptr(self.ws), self.ws.numel(), C.byref(n),
                                             dv.stream_ptr())
            if rc == -2 and n.value > self.max_entries:  # BF_E_WORKSPACE: grow to the exact count
                self.max_entries = n.value
                continue
            _lib.check(rc, "bf_shell_bin_halos")
            self.nentries = n.value
            return self
        raise _lib.BFError("bf_shell_bin_halos: workspace retry failed")

    # ---- stage 2: pair evaluation ----------------------------------------------------------------
    def accumulate(self, out=None, add=False):
        """Returns the device accumulator of the engine's pixel range: map [n] or pix_offsets [n, 3].  ``add``: sum on
        top of what ``out`` holds (gather variant; halo chunks of one catalogue evaluated one after the other)."""
        n = self.pix_last - self.pix_first
        shape = (n,) if self.mode == PAINT else (n, 3)
        self.cnt.zero_()
        if self.variant == "gather":
            if out is None:
                out = dv.empty(shape, dev=self.dev)
            _lib.call("bf_shell_gather", self.nside, self.mode, C.byref(self.desc), C.byref(self.table.desc),
                      self.bands[0], self.bands[1], dv.ptr(self.ws), dv.ptr(out), self.pix_first, 1 if add else 0,
                      dv.ptr(self.cnt), dv.stream_ptr())
        else:
            if add:
                raise _lib.BFError("chunked accumulation is a feature of the gather variant")
            if self.bands != (0, self.lib.bf_shell_num_bands(self.nside)):
                raise _lib.BFError("the atomic variant does not shard by ring band")
            if out is None:
                out = dv.zeros(shape, dev=self.dev)
            else:
                out.zero_()
            fn = "bf_shell_paint_atomic" if self.mode == PAINT else "bf_shell_baryonify_atomic"
            _lib.call(fn, self.nside, C.byref(self.desc), C.byref(self.table.desc), dv.ptr(out), dv.ptr(self.cnt),
                      dv.stream_ptr())
        return out

    def pair_count(self):
        self.npairs = int(self.cnt.item())
        return self.npairs

    # ---- stage 3 (baryonify): re-binning ------------------------------------------------------------
    def regrid(self, orig_d, off, new_d=None, method="gather"):
        """orig_d / off are the engine's pixel slice; new_d is the full sphere.  "gather" = deterministic destination
        gather (whole sphere only; a pixel slice, or a displacement window that is too wide, takes "sorted");
        "sorted" = deterministic counting sort by destination + in-order segmented sum; both give the reference's
        summation order and the same bits.  "atomic" = fp64 RED.ADD.  ``self.regrid_used`` names what ran."""
        if method not in ("gather", "sorted", "atomic"):
            raise ValueError(f"unknown regrid method {method!r}")
        whole = self.pix_first == 0 and self.pix_last == self.npix
        fresh = new_d is None
        if fresh:
            new_d = dv.empty(self.npix, dev=self.dev) if (method == "gather" and whole) else dv.zeros(self.npix, dev=self.dev)
        elif not (method == "gather" and whole):  # (the whole-sphere destination gather writes every pixel itself)
            new_d.zero_()
        if method == "gather" and whole:
            need = self.lib.bf_shell_regrid_gather_ws_bytes(self.nside)
            if getattr(self, "_rg_ws", None) is None or self._rg_ws.numel() < need:
                self._rg_ws = None
                self._rg_ws = torch.empty(need, dtype=torch.uint8, device=self.dev)
            win = (C.c_int * 3)()
            rc = self.lib.bf_shell_regrid_gather(self.nside, dv.ptr(orig_d), dv.ptr(off), dv.ptr(new_d),
                                                 dv.ptr(self._rg_ws), self._rg_ws.numel(), win, dv.stream_ptr())
            self.regrid_window = tuple(win)
            if rc != -3:  # BF_E_WINDOW: displacements span too many pixels for the gather -> counting sort
                _lib.check(rc, "bf_shell_regrid_gather")
                self.regrid_used = "gather"
                return new_d
            new_d.zero_()
        if method == "gather" and not whole and self.pix_last > self.pix_first:
            # ring-band shard: PARTIAL sums of the shard's own sources over the destinations they can reach; the shards
            # then exchange the boundary mass (sharding.exchange_boundary)
            need = self.lib.bf_shell_regrid_gather_ws_bytes(self.nside)
            if getattr(self, "_rg_ws", None) is None or self._rg_ws.numel() < need:
                self._rg_ws = None
                self._rg_ws = torch.empty(need, dtype=torch.uint8, device=self.dev)
            win = (C.c_int * 3)()
            rc = self.lib.bf_shell_regrid_gather_emit_range(self.nside, self.pix_first, self.pix_last, dv.ptr(orig_d),
                                                            dv.ptr(off), dv.ptr(self._rg_ws), self._rg_ws.numel(), win,
                                                            dv.stream_ptr())
            self.regrid_window = tuple(win)
            if rc != -3:
                _lib.check(rc, "bf_shell_regrid_gather_emit_range")
                reach = (max(win[1], -win[0]) + 1) * 4 * self.nside
                _lib.call("bf_shell_regrid_gather_range", self.nside, dv.ptr(self._rg_ws), self._rg_ws.numel(), win,
                          max(0, self.pix_first - reach), min(self.npix, self.pix_last + reach), dv.ptr(new_d),
                          dv.stream_ptr())
                self.regrid_used = "gather"
                return new_d
        if method in ("gather", "sorted"):
            need = self.lib.bf_shell_regrid_sorted_ws_bytes(self.nside, self.pix_last - self.pix_first)
            if getattr(self, "_rg_ws", None) is None or self._rg_ws.numel() < need:
                self._rg_ws = None
                self._rg_ws = torch.empty(need, dtype=torch.uint8, device=self.dev)
            _lib.call("bf_shell_regrid_sorted", self.nside, self.pix_first, self.pix_last, dv.ptr(orig_d), dv.ptr(off),
                      dv.ptr(new_d), dv.ptr(self._rg_ws), self._rg_ws.numel(), dv.stream_ptr())
            self.regrid_used = "sorted"
        else:
            _lib.call("bf_shell_regrid_range", self.nside, self.pix_first, self.pix_last, dv.ptr(orig_d), dv.ptr(off),
                      dv.ptr(new_d), dv.stream_ptr())
            self.regrid_used = "atomic"
        return new_d

    def regrid_shard(self, orig_d, off, F, lo_neighbour, hi_neighbour, method="gather"):
        """Re-binning of a ring-band shard into a LOCAL array over [pix_first - F, pix_last + F) (a side without a
        neighbour has no margin): the shard's partial sums, own pixels complete up to the neighbours' boundary mass
        (``sharding.exchange_margins`` moves and adds it).  F = ``sharding.margin_pixels``, which the shard form of the
        destination gather never exceeds (it answers BF_E_WINDOW first); then -- or for another ``method`` -- the
        counting sort / atomics run into a whole-sphere array and the window is cut out of it.
        Returns (local tensor, device flag or None: mass landed outside the window -- the caller raises on it AFTER the
        exchange, so that no rank is left waiting)."""
        w0 = self.pix_first - (F if lo_neighbour else 0)
        w1 = self.pix_last + (F if hi_neighbour else 0)
        assert 0 <= w0 and w1 <= self.npix
        if method == "gather" and self.pix_last > self.pix_first:
            need = self.lib.bf_shell_regrid_gather_ws_bytes(self.nside)
            if getattr(self, "_rg_ws", None) is None or self._rg_ws.numel() < need:
                self._rg_ws = None
                self._rg_ws = torch.empty(need, dtype=torch.uint8, device=self.dev)
            win = (C.c_int * 3)()
            rc = self.lib.bf_shell_regrid_gather_emit_range(self.nside, self.pix_first, self.pix_last, dv.ptr(orig_d),
                                                            dv.ptr(off), dv.ptr(self._rg_ws), self._rg_ws.numel(), win,
                                                            dv.stream_ptr())
            self.regrid_window = tuple(win)
            if rc != -3:
                _lib.check(rc, "bf_shell_regrid_gather_emit_range")
                reach = (max(win[1], -win[0]) + 1) * 4 * self.nside  # <= F
                d0, d1 = max(w0, self.pix_first - reach), min(w1, self.pix_last + reach)
                local = self._local_buf(w1 - w0)
                if d0 > w0:
                    local[:d0 - w0].zero_()
                if d1 < w1:
                    local[d1 - w0:].zero_()
                base = C.c_void_p(local.data_ptr() - 8 * w0)  # the kernel indexes with sphere pixel numbers
                _lib.call("bf_shell_regrid_gather_range", self.nside, dv.ptr(self._rg_ws), self._rg_ws.numel(), win,
                          d0, d1, base, dv.stream_ptr())
                self.regrid_used = "gather"
                return local, None
            method = "sorted"
        full = self.regrid(orig_d, off, method="sorted" if method == "gather" else method)
        stray = (full[:w0] != 0).any() | (full[w1:] != 0).any()
        return full[w0:w1].clone(), stray

    def _local_buf(self, n):
        buf = getattr(self, "_rg_local", None)
        if buf is None or buf.numel() != n:
            self._rg_local = buf = torch.empty(n, dtype=torch.float64, device=self.dev)
        return buf

    def regrid_to_host(self, orig_d, off, method="gather", pinned=True, chunks=8):
        """Re-binning + read-back of the new map -> (host array, device scalar sum(new map), seconds of the device part).

        With the destination-gather variant the map is PRODUCED in destination ranges, so every finished range starts its
        way to the host (side stream, page-locked target) while the next one is computed; the other variants (and a
        window that is too wide) re-bin first and copy afterwards."""
        whole = self.pix_first == 0 and self.pix_last == self.npix
        if method == "gather" and whole and pinned and chunks > 1:
            need = self.lib.bf_shell_regrid_gather_ws_bytes(self.nside)
            if getattr(self, "_rg_ws", None) is None or self._rg_ws.numel() < need:
                self._rg_ws = None
                self._rg_ws = torch.empty(need, dtype=torch.uint8, device=self.dev)
            main = torch.cuda.current_stream(self.dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(main)
            win = (C.c_int * 3)()
            rc = self.lib.bf_shell_regrid_gather_emit(self.nside, dv.ptr(orig_d), dv.ptr(off), dv.ptr(self._rg_ws),
                                                      self._rg_ws.numel(), win, dv.stream_ptr())
            self.regrid_window = tuple(win)
            if rc == 0:
                new_d = dv.empty(self.npix, dev=self.dev)  # (every destination is written by the gather)
                out = torch.empty(self.npix, dtype=torch.float64, pin_memory=True)
                side = torch.cuda.Stream(device=self.dev)
                edges = [self.npix * i // chunks for i in range(chunks + 1)]
                for a, b in zip(edges[:-1], edges[1:]):
                    _lib.call("bf_shell_regrid_gather_range", self.nside, dv.ptr(self._rg_ws), self._rg_ws.numel(), win,
                              a, b, dv.ptr(new_d), dv.stream_ptr())
                    done = torch.cuda.Event()
                    done.record(main)
                    side.wait_event(done)
                    with torch.cuda.stream(side):
                        out[a:b].copy_(new_d[a:b], non_blocking=True)
                e1.record(main)
                new_sum = new_d.sum()
                new_d.record_stream(side)
                side.synchronize()
                main.synchronize()
                self.regrid_used = "gather"
                return out.numpy(), new_sum, e0.elapsed_time(e1) * 1e-3
            if rc != -3:  # (BF_E_WINDOW: displacements span too many pixels for the gather -> counting sort)
                _lib.check(rc, "bf_shell_regrid_gather_emit")
            method = "sorted"
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        new_d = self.regrid(orig_d, off, method=method)
        e1.record()
        new_sum = new_d.sum()
        host = dv.to_host(new_d, pinned)
        return host, new_sum, e0.elapsed_time(e1) * 1e-3

    def baryonify_to_host(self, upload, ngroups=8, slack_rings=8):
        """BaryonifyShell end to end on the device side, pipelined by ring-band groups (whole sphere, gather variants):

            for each group g (ascending rings):  pair stage of g  ->  re-binning deposits of g's pixels (emit)
                                                 ->  destinations all of whose sources are emitted: gather + read-back

        so that the 1.6 GB read-back (and the map upload: ``upload`` is a ``ChunkedUpload`` cut at the same pixel edges)
        run under the pair kernels of the later groups instead of after them.  A destination of ring r can only receive
        from source rings r - dr_max .. r - dr_min; rings up to (last emitted ring) - |dr_min| - ``slack_rings`` are
        finalised after every group, and the assumption that no LATER source reaches back further than that is checked
        at the end against the final window (else the re-binning is simply redone in one piece).  Same deposits, same
        ascending-source order, hence the same bits as ``regrid(method="gather")``.
        Returns (host map, device scalar sum(new map), device scalar sum(old map))."""
        assert self.variant == "gather" and self.mode == BARYONIFY
        assert self.pix_first == 0 and self.pix_last == self.npix
        ns, npix, lib = self.nside, self.npix, self.lib
        groups, edges = band_groups(ns, ngroups)
        assert edges == upload.edges, "the map upload must be cut at the group edges"
        need = lib.bf_shell_regrid_gather_ws_bytes(ns)
        if getattr(self, "_rg_ws", None) is None or self._rg_ws.numel() < need:
            self._rg_ws = None
            self._rg_ws = torch.empty(need, dtype=torch.uint8, device=self.dev)
        off = dv.empty((npix, 3), dev=self.dev)
        new_d = dv.empty(npix, dev=self.dev)
        out = torch.empty(npix, dtype=torch.float64, pin_memory=True)
        main, side = torch.cuda.current_stream(self.dev), torch.cuda.Stream(device=self.dev)
        orig_d = upload.out
        self.cnt.zero_()
        last_ring = 4 * ns - 1
        win_u = None          # union of the groups' windows so far: [dr_min, dr_max, dj_max]
        done_pix, done_ring, min_margin = 0, 0, None
        wide = False          # a group's window too wide for the gather: finish the pair stage, re-bin in one piece
        t_emit = 0.0

        def c_win(w):
            return (C.c_int * 3)(*(w if w is not None else (1, 0, 0)))

        def finalise(p1):
            nonlocal done_pix
            _lib.call("bf_shell_regrid_gather_range", ns, dv.ptr(self._rg_ws), self._rg_ws.numel(), c_win(win_u),
                      done_pix, p1, dv.ptr(new_d), dv.stream_ptr())
            done = torch.cuda.Event()
            done.record(main)
            side.wait_event(done)
            with torch.cuda.stream(side):
                out[done_pix:p1].copy_(new_d[done_pix:p1], non_blocking=True)
            done_pix = p1

        for gi, (lo, hi) in enumerate(groups):
            _lib.call("bf_shell_gather", ns, self.mode, C.byref(self.desc), C.byref(self.table.desc), lo, hi,
                      dv.ptr(self.ws), dv.ptr(off), 0, 0, dv.ptr(self.cnt), dv.stream_ptr())
            if wide:
                continue
            a, b = edges[gi], edges[gi + 1]
            upload.wait(gi)
            win = (C.c_int * 3)()
            rc = lib.bf_shell_regrid_gather_emit_part(ns, a, b, dv.ptr(orig_d[a:b]), dv.ptr(off[a:b]), dv.ptr(self._rg_ws),
                                                      self._rg_ws.numel(), win, dv.stream_ptr())
            if rc == -3:  # BF_E_WINDOW
                wide = True
                continue
            _lib.check(rc, "bf_shell_regrid_gather_emit_part")
            if win[1] >= win[0]:
                win_u = list(win) if win_u is None else [min(win_u[0], win[0]), max(win_u[1], win[1]), max(win_u[2], win[2])]
                cost = (win_u[1] - win_u[0] + 1) * (2 * win_u[2] + 1)
                if cost > 165:  # BF_SHELL_GATHER_MAX_WINDOW for the union
                    wide = True
                    continue
            emitted = min(8 * hi, last_ring)
            if gi == len(groups) - 1:
                finalise(npix)
            else:
                margin = max(0, -(win_u[0] if win_u is not None else 0)) + int(slack_rings)
                r_fin = emitted - margin
                if r_fin > done_ring:
                    min_margin = margin if min_margin is None else min(min_margin, margin)
                    finalise(ring_first_pixel(ns, r_fin + 1))
                    done_ring = r_fin
        old_sum = upload.result().sum()
        self.pair_count()  # (synchronises the main stream)
        redo = wide or (win_u is not None and min_margin is not None and -win_u[0] > min_margin)
        if redo:  # rare: one group's deposits reached back further than assumed, or a very wide window
            side.synchronize()
            new_h, new_sum, _ = self.regrid_to_host(orig_d, off, method="gather")
            self.regrid_pipelined = False
            return new_h, new_sum, old_sum
        self.regrid_used, self.regrid_pipelined = "gather", True
        self.regrid_window = tuple(win_u) if win_u is not None else (0, 0, 0)
        new_sum = new_d.sum()
        new_d.record_stream(side)
        side.synchronize()
        main.synchronize()
        return out.numpy(), new_sum, old_sum

    def accumulate_to_host(self, chunks=8, out_h=None):
        """Pair stage + read-back of the accumulator (PaintProfilesShell) -> host array.  Pixel owners finish ring band
        by ring band, so the gather kernel is launched over ``chunks`` groups of the engine's bands and every finished
        group is copied to page-locked host memory on a side stream while the next one is evaluated.  ``out_h``: a
        page-locked host tensor for the engine's pixel range (a ring-band shard writes its piece of the shared result
        array); default: a fresh one from torch's caching host allocator."""
        nloc = self.pix_last - self.pix_first
        shape = (nloc,) if self.mode == PAINT else (nloc, 3)
        if self.variant != "gather" or chunks < 2 or self.nh == 0 or self.bands[1] - self.bands[0] < 2:
            out = self.accumulate()
            self.pair_count()
            if out_h is None:
                return dv.to_host(out, True)
            out_h.copy_(out.view(out_h.shape), non_blocking=True)
            torch.cuda.current_stream(self.dev).synchronize()
            return out_h.numpy()
        from .sharding import band_ends, band_pixel_range
        lo0, hi0 = self.bands
        k = min(chunks, hi0 - lo0)
        ends = band_ends(self.nside)[lo0:hi0 + 1]  # first pixel after bands .. b - 1, for b = lo0 .. hi0
        cuts = [lo0] + [lo0 + int(np.searchsorted(ends, self.pix_first + nloc * i // k)) for i in range(1, k)] + [hi0]
        groups = [(a, b) for a, b in zip(cuts[:-1], cuts[1:]) if b > a]  # near-equal pixel counts
        out_d = dv.empty(shape, dev=self.dev)
        if out_h is None:
            out_h = torch.empty(shape, dtype=torch.float64, pin_memory=True)
        main, side = torch.cuda.current_stream(self.dev), torch.cuda.Stream(device=self.dev)
        self.cnt.zero_()
        for lo, hi in groups:
            _lib.call("bf_shell_gather", self.nside, self.mode, C.byref(self.desc), C.byref(self.table.desc), lo, hi,
                      dv.ptr(self.ws), dv.ptr(out_d), self.pix_first, 0, dv.ptr(self.cnt), dv.stream_ptr())
            a, b = band_pixel_range(self.nside, lo, hi)
            a, b = a - self.pix_first, b - self.pix_first
            done = torch.cuda.Event()
            done.record(main)
            side.wait_event(done)
            with torch.cuda.stream(side):
                out_h[a:b].copy_(out_d[a:b], non_blocking=True)
        out_d.record_stream(side)
        self.pair_count()  # (synchronises the main stream)
        side.synchronize()
        return out_h.numpy()
