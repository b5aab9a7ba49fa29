"""Periodic 2-D / 3-D grid runners on the GPU: ``BaryonifyGrid`` and ``PaintProfilesGrid``.

Drop-in for ``BaryonForge/Runners/Map2DRunner.py`` (``DefaultRunnerGrid`` :171-366, ``BaryonifyGrid``
:370-615, ``PaintProfilesGrid`` :618-814): same constructor and ``process()`` contract.
"""
from __future__ import annotations

import ctypes as C
import time

import numpy as np
import torch

from .. import _device as dv
from .. import _lib, host
from .._ccl import default_mass_def
from ..tables import displacement_host_table, profile_host_table

__all__ = ["DefaultRunnerGrid", "BaryonifyGrid", "PaintProfilesGrid"]

GRID_BARYONIFY, GRID_PAINT = 0, 1  # BF_GRID_MODE_* of the C ABI (the shell entries number them the other way round)


class DefaultRunnerGrid:
    def __init__(self, HaloNDCatalog, GriddedMap, epsilon_max, model, use_ellipticity=False, mass_def=None,
                 verbose=True, *, device=None, regrid="gather", variant="gather"):
        self.HaloNDCatalog = HaloNDCatalog
        self.GriddedMap = GriddedMap
        self.cosmo = HaloNDCatalog.cosmology
        self.model = model
        self.epsilon_max = epsilon_max
        self.mass_def = mass_def if mass_def is not None else default_mass_def()
        self.verbose = verbose
        self.use_ellipticity = use_ellipticity
        self.device = device
        # re-binning: "gather" (deterministic, default; falls back to "sorted" for very wide displacement windows)
        # | "sorted" (deterministic, 25 B of workspace per deposit) | "atomic"
        self.regrid = regrid
        # pair loop: "gather" (cell-owner, deterministic catalogue-order sums, default) | "atomic" (halo-centric)
        self.variant = variant
        self.rows = None  # (a_first, a_last): slab of the first grid axis owned by this rank (sharding.ShardedGrid)
        self.npairs = None
        self.timings = {}
        if use_ellipticity:  # Map2DRunner.py:266-272
            names = HaloNDCatalog.cat.dtype.names
            assert "q_ell" in names, "The 'q_ell' column is missing, but you set use_ellipticity = True"
            if not GriddedMap.is2D:
                assert "c_ell" in names, "The 'c_ell' column is missing, but you set use_ellipticity = True"
            assert "A_ell" in names, "The 'A_ell' column is missing, but you set use_ellipticity = True"

    def _progress(self, what):
        """Per-stage progress lines under the reference's ``verbose`` flag (its tqdm bar over halos, Map2DRunner.py:476 /
        :714, has no per-halo counterpart here: the loop is a handful of kernel launches)."""
        if self.verbose:
            import sys
            print("[%s] %s" % (type(self).__name__, what), file=sys.stderr, flush=True)

    def _prepare(self, table_host, baryonify, dev):
        self._progress("host set-up of %d halos, table upload, (z, M) collapse" % len(self.HaloNDCatalog.cat))
        keys = vars(self.model).get("p_keys", [])
        if len(keys) > 0:  # Map2DRunner.py:469-474 / :707-712
            txt = (f"You asked to use {keys} properties in Baryonification. You must pass a ParamTabulatedProfile"
                   f"as the model. You have passed {type(self.model)} instead")
            tab = getattr(self.model, "raw_input_2D", None)
            assert (not baryonify) and tab is not None and np.ndim(tab) == 3 + len(keys), txt
        G = self.GriddedMap
        rmat = host.grid_rmat(self.HaloNDCatalog, G, self.use_ellipticity, baryonify)  # 3-D raises (:565 / :790)
        table = dv.DeviceTable(table_host, dev=dev)
        s = host.grid_scalars(self.HaloNDCatalog, G, self.mass_def, self.epsilon_max, self.model, baryonify)
        t = {k: dv.to_dev(s[k], dev=dev) for k in ("pos", "cut", "lnr_shift", "z_in", "M_in")}
        t["Nsize"] = dv.to_dev(s["Nsize"], dtype=np.int64, dev=dev)
        extras = host.property_columns(self.HaloNDCatalog.cat, self.model)
        t["extras"] = dv.to_dev(extras, dev=dev) if extras is not None else None
        rows = table.collapse(t["z_in"], t["M_in"], t["extras"])
        t["Rmat"] = dv.to_dev(rmat, dev=dev) if rmat is not None else None
        desc = _lib.GridHalos(t["pos"].data_ptr(), t["Nsize"].data_ptr(), t["cut"].data_ptr(),
                              t["lnr_shift"].data_ptr(), rows.data_ptr(),
                              t["Rmat"].data_ptr() if rmat is not None else None, s["cut"].size)
        bins = dv.to_dev(G.bins, dev=dev)
        return table, s, desc, bins, (t, rows)

    def _accumulate(self, mode, s, desc, bins, table, out, cnt, dev, row_runs=None, after_run=None):
        """The per-halo loop (Map2DRunner.py:476-580 / :714-810) into ``out``: cell-owner gather, or atomics.
        ``row_runs``: evaluate the cell owners in these runs of rows of the first axis, one launch each, and call
        ``after_run(a, b)`` after every launch (the caller starts the read-back of the finished rows there)."""
        G = self.GriddedMap
        lib = _lib.load(require_cuda=True)
        dim, nh = s["dim"], s["cut"].size
        self._progress("halo binning into grid tiles + pair stage (%s)" % self.variant)
        self.variant_used = self.variant
        if self.variant == "gather" and nh > 0 and int(np.max(s["Nsize"])) > G.Npix:
            self.variant_used = "atomic"  # a cut-out wider than the grid: numpy's duplicate-dropping += semantics
        if self.variant_used == "gather" and nh > 0:
            max_entries, n = 0, C.c_int64()
            for _ in range(2):  # counting pass first (exact size), then the real one
                nb = lib.bf_grid_bin_ws_bytes(dim, G.Npix, nh, max_entries)
                ws = torch.empty(nb, dtype=torch.uint8, device=dev)
                rc = lib.bf_grid_bin_halos(dim, G.Npix, dv.ptr(bins), C.byref(desc), table.nr, max_entries, dv.ptr(ws),
                                           nb, C.byref(n), dv.stream_ptr())
                if rc == -2 and n.value > max_entries:
                    max_entries = n.value
                    continue
                _lib.check(rc, "bf_grid_bin_halos")
                break
            self.nentries = n.value
            runs = [self.rows if self.rows is not None else (0, G.Npix)]
            if row_runs and self.rows is None:
                runs = row_runs
            for a0, a1 in runs:
                _lib.call("bf_grid_gather_rows", dim, mode, G.Npix, float(G.res), C.byref(desc), C.byref(table.desc),
                          max_entries, dv.ptr(ws), dv.ptr(out), a0, a1, dv.ptr(cnt), dv.stream_ptr())
                if after_run is not None and runs is row_runs:
                    after_run(a0, a1)
        elif nh > 0:
            if self.rows is not None:
                raise _lib.BFError("slab-sharded grids need the cell-owner gather (variant='gather')")
            _lib.call("bf_grid_paint_atomic" if mode else "bf_grid_baryonify_atomic", dim, G.Npix, dv.ptr(bins),
                      float(G.res), C.byref(desc), C.byref(table.desc), dv.ptr(out), dv.ptr(cnt), dv.stream_ptr())
        self.npairs = int(cnt.item())


class BaryonifyGrid(DefaultRunnerGrid):
    def offsets_device(self, dev):
        G = self.GriddedMap
        t0 = time.perf_counter()
        table, s, desc, bins, keep = self._prepare(displacement_host_table(self.model), True, dev)
        off = dv.zeros((G.map.size, s["dim"]), dev=dev)
        cnt = dv.pair_counter(dev)
        torch.cuda.synchronize(dev)
        self.timings["setup"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        self._accumulate(GRID_BARYONIFY, s, desc, bins, table, off, cnt, dev)
        self.timings["accumulate"] = time.perf_counter() - t0
        del keep
        return off

    def process(self):
        dev = dv.device(self.device)
        G = self.GriddedMap
        with torch.cuda.device(dev):
            t00 = time.perf_counter()
            orig = np.ascontiguousarray(G.map, dtype=np.float64).reshape(-1)
            up = dv.AsyncUpload(orig, dev)  # the map travels while the halos are set up and evaluated
            off = self.offsets_device(dev)
            t0 = time.perf_counter()
            orig_d = up.result()
            new_d = dv.zeros(orig.size, dev=dev)
            dim = 2 if G.is2D else 3
            self._progress("re-binning (%s) + read-back" % self.regrid)
            lib = _lib.load(require_cuda=True)
            mode = self.regrid
            new_map = None
            if mode == "gather":
                # destination gather, produced in runs of rows of the first axis: every finished run is read back (side
                # stream, page-locked target) while the next one is computed
                ws = torch.empty(lib.bf_grid_regrid_gather_ws_bytes(), dtype=torch.uint8, device=dev)
                lohi, win = (C.c_int * 6)(), (C.c_int * 3)()
                _lib.call("bf_grid_regrid_window", dim, G.Npix, dv.ptr(off), 0, G.Npix, dv.ptr(ws), ws.numel(), lohi,
                          dv.stream_ptr())
                rowlen = orig.size // G.Npix
                nchunk = 8 if G.Npix >= 64 else 1
                out_h = torch.empty(orig.size, dtype=torch.float64, pin_memory=True)
                main, side = torch.cuda.current_stream(dev), torch.cuda.Stream(device=dev)
                for i in range(nchunk):
                    a, b = G.Npix * i // nchunk, G.Npix * (i + 1) // nchunk
                    rc = lib.bf_grid_regrid_gather_rows(dim, G.Npix, dv.ptr(off), dv.ptr(orig_d), dv.ptr(new_d), lohi,
                                                        a, b, win, dv.stream_ptr())
                    self.regrid_window = tuple(win)
                    if rc == -3:  # BF_E_WINDOW: displacements span too many cells for the gather (decided on the window,
                        mode = "sorted"  # i.e. at the first run: nothing has been written yet)
                        break
                    _lib.check(rc, "bf_grid_regrid_gather_rows")
                    done = torch.cuda.Event()
                    done.record(main)
                    side.wait_event(done)
                    with torch.cuda.stream(side):
                        out_h[a * rowlen:b * rowlen].copy_(new_d[a * rowlen:b * rowlen], non_blocking=True)
                if mode == "gather":
                    new_d.record_stream(side)
                    new_map = out_h
            self.regrid_used = mode
            if mode == "gather":
                pass
            elif mode == "sorted":
                nb = lib.bf_grid_regrid_sorted_ws_bytes(dim, G.Npix)
                ws = torch.empty(nb, dtype=torch.uint8, device=dev)
                _lib.call("bf_grid_regrid_sorted", dim, G.Npix, dv.ptr(off), dv.ptr(orig_d), dv.ptr(new_d), dv.ptr(ws),
                          nb, dv.stream_ptr())
                del ws
            else:
                _lib.call("bf_grid_regrid", dim, G.Npix, dv.ptr(off), dv.ptr(orig_d), dv.ptr(new_d), dv.stream_ptr())
            new_sum_d, old_sum_d = new_d.sum(), orig_d.sum()
            torch.cuda.synchronize(dev)
            self.timings["regrid"] = time.perf_counter() - t0
            t0 = time.perf_counter()
            if new_map is not None:
                side.synchronize()
                new_map = new_map.numpy().reshape(G.map.shape)
            else:
                new_map = dv.to_host(new_d, True).reshape(G.map.shape)
            new_sum, old_sum = float(new_sum_d.item()), float(old_sum_d.item())  # Map2DRunner.py:611-613
            self.timings["d2h"] = time.perf_counter() - t0
            self.timings["process"] = time.perf_counter() - t00
        assert np.isclose(new_sum, old_sum), \
            "ERROR in pixel regridding, sum(new_map) [%0.14e] != sum(oldmap) [%0.14e]" % (new_sum, old_sum)
        return new_map


class PaintProfilesGrid(DefaultRunnerGrid):
    def paint_device(self, dev, row_runs=None, after_run=None):
        """The painted map as a device tensor [Npix^D] (rows outside ``self.rows`` stay zero)."""
        G = self.GriddedMap
        table, s, desc, bins, keep = self._prepare(profile_host_table(self.model, "2D" if G.is2D else "3D"), False, dev)
        out = dv.zeros(G.map.size, dev=dev)
        cnt = dv.pair_counter(dev)
        if after_run is not None:
            after_run(None, None, out)  # hand the accumulator to the caller's read-back
        self._accumulate(GRID_PAINT, s, desc, bins, table, out, cnt, dev, row_runs, after_run)
        del keep
        return out

    def process(self):
        dev = dv.device(self.device)
        G = self.GriddedMap
        with torch.cuda.device(dev):
            t0 = time.perf_counter()
            n = G.Npix
            if self.variant == "gather" and self.rows is None and n >= 128 and G.map.size >= (1 << 23):  # >= 64 MB
                # cell owners finish run by run of rows (tile-aligned): a finished run's read-back overlaps the next launch
                rowlen = G.map.size // n
                edges = sorted(set([0, n] + [min(n, (n * i // 8 + 15) // 16 * 16) for i in range(1, 8)]))
                runs = [(a, b) for a, b in zip(edges[:-1], edges[1:]) if b > a]
                out_h = torch.empty(G.map.size, dtype=torch.float64, pin_memory=True)
                main, side = torch.cuda.current_stream(dev), torch.cuda.Stream(device=dev)
                copied, box = [], {}

                def after_run(a, b, acc=None):  # (no reference to itself: a cycle would keep the pinned block alive)
                    if acc is not None:
                        box["out"] = acc
                        return
                    done = torch.cuda.Event()
                    done.record(main)
                    side.wait_event(done)
                    with torch.cuda.stream(side):
                        out_h[a * rowlen:b * rowlen].copy_(box["out"][a * rowlen:b * rowlen], non_blocking=True)
                    copied.append((a, b))

                out = self.paint_device(dev, runs, after_run)
                box.clear()
                self.timings["accumulate"] = time.perf_counter() - t0
                if copied:  # (the atomic fall-back of a too-wide cut-out evaluates everything at once)
                    out.record_stream(side)
                    side.synchronize()
                    res = out_h.numpy().reshape(G.map.shape)
                else:
                    res = dv.to_host(out, True).reshape(G.map.shape)
                self.timings["process"] = time.perf_counter() - t0
                return res
            out = self.paint_device(dev)
            self.timings["accumulate"] = time.perf_counter() - t0
            res = dv.to_host(out, True).reshape(G.map.shape)
            self.timings["process"] = time.perf_counter() - t0
        return res
