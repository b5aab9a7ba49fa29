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
        self.nh = int(np.asarray(scalars["a"]).size)
        self.table = None
        self.ws = None
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

    def recollapse(self):
        self.rows = self.table.collapse(self.t["z_in"], self.t["M_in"], self.t["extras"])
        self.desc.rows = self.rows.data_ptr()

    # ---- stage 1: halo binning into sky cells (gather variant) ----------------------------------
    def bin(self):
        if self.variant != "gather":
            return self
        # first call: a counting pass with a key-less workspace returns the exact number of (cell, halo) entries
        # (BF_E_WORKSPACE + *nentries_host), the second call is sized for it; later calls reuse the workspace
        if self.ws is None:
            self.max_entries = 0
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
        if new_d is None:
            new_d = dv.zeros(self.npix, dev=self.dev)
        else:
            new_d.zero_()
        if method not in ("gather", "sorted", "atomic"):
            raise ValueError(f"unknown regrid method {method!r}")
        whole = self.pix_first == 0 and self.pix_last == self.npix
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
                new_d = dv.zeros(self.npix, dev=self.dev)
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

    def accumulate_to_host(self, chunks=8):
        """Pair stage + read-back of the accumulator (PaintProfilesShell) -> host array.  Pixel owners finish ring band
        by ring band, so the gather kernel is launched over ``chunks`` groups of bands (equal pixel counts) and every
        finished group is copied to page-locked host memory on a side stream while the next one is evaluated."""
        whole = self.pix_first == 0 and self.pix_last == self.npix
        if self.variant != "gather" or not whole or chunks < 2 or self.nh == 0:
            out = self.accumulate()
            self.pair_count()
            return dv.to_host(out, True)
        from .sharding import band_pixel_range, band_split
        nb = self.lib.bf_shell_num_bands(self.nside)
        groups = [g for g in band_split(self.nside, min(chunks, nb)) if g[1] > g[0]]
        shape = (self.npix,) if self.mode == PAINT else (self.npix, 3)
        out_d = dv.empty(shape, dev=self.dev)
        out_h = torch.empty(shape, dtype=torch.float64, pin_memory=True)
        main, side = torch.cuda.current_stream(self.dev), torch.cuda.Stream(device=self.dev)
        self.cnt.zero_()
        for lo, hi in groups:
            _lib.call("bf_shell_gather", self.nside, self.mode, C.byref(self.desc), C.byref(self.table.desc), lo, hi,
                      dv.ptr(self.ws), dv.ptr(out_d), 0, 0, dv.ptr(self.cnt), dv.stream_ptr())
            a, b = band_pixel_range(self.nside, lo, hi)
            done = torch.cuda.Event()
            done.record(main)
            side.wait_event(done)
            with torch.cuda.stream(side):
                out_h[a:b].copy_(out_d[a:b], non_blocking=True)
        out_d.record_stream(side)
        self.pair_count()  # (synchronises the main stream)
        side.synchronize()
        return out_h.numpy()
