"""HEALPix shell runners on the GPU: ``BaryonifyShell`` and ``PaintProfilesShell``.

Drop-in for ``BaryonForge/Runners/HealpixRunner.py`` (``DefaultRunner`` :74-220, ``BaryonifyShell`` :223-349,
``PaintProfilesShell`` :352-447): same constructor (positional order included -- ``SplitJoinParallel``
re-instantiates positionally, ``utils/Parallelize.py:271``), same public attributes, ``process()``
returns a float64 numpy map.  The serial per-halo loop is replaced by CUDA kernels behind the C ABI.
"""
from __future__ import annotations

import ctypes as C
import time

import numpy as np
import torch

from .. import _device as dv
from .. import _lib, engine, host
from .._ccl import default_mass_def
from ..tables import displacement_host_table, profile_host_table

__all__ = ["DefaultRunner", "BaryonifyShell", "PaintProfilesShell"]


class DefaultRunner:
    def __init__(self, HaloLightConeCatalog, LightconeShell, epsilon_max, model, use_ellipticity=False,
                 mass_def=None, verbose=True, *, device=None, variant="gather", regrid="gather", pinned_output=True):
        self.HaloLightConeCatalog = HaloLightConeCatalog
        self.LightconeShell = LightconeShell
        self.cosmo = HaloLightConeCatalog.cosmology
        self.model = model
        self.epsilon_max = epsilon_max
        self.mass_def = mass_def if mass_def is not None else default_mass_def()
        self.verbose = verbose
        self.use_ellipticity = use_ellipticity
        self.device = device
        self.variant = variant          # pair stage: "gather" (deterministic, default) | "atomic"
        self.regrid = regrid            # re-binning: "gather" (deterministic, default; falls back to "sorted") | "sorted" | "atomic"
        self.pinned_output = pinned_output
        # opt-in (pipeline_chunks > 1): catalogues of at least `pipeline_min_halos` are cut into runs of consecutive halos
        # and the host set-up of run k+1 overlaps the GPU stages of run k (sums continue on the device in catalogue
        # order).  Off by default: on the 16-vCPU hosts of this pool the host threads, the pinned-memory copies and the
        # launching thread compete for the same cores and the end-to-end time does not improve (scripts/e2e_shell_pipeline.py)
        self.pipeline_chunks, self.pipeline_min_halos = 1, 400_000
        self.pipeline_upload_after = 0  # index of the halo run after whose host set-up the map upload starts
        self.npairs = None
        self.timings = {}
        if use_ellipticity:  # HealpixRunner.py:164-165
            raise NotImplementedError("You have set use_ellipticity = True, but this not yet implemented for HealpixRunner")

    # opt-in (lower SETUP_RUNS_MIN_HALOS): measured on the 16-vCPU hosts of this pool the streamed set-up LOSES (process()
    # 111 ms with the size hint alone, 114 / 129 / 155 ms with 2 / 4 / 8 runs: the enqueueing thread and the host threads
    # queue for the interpreter lock), so it is off by default
    SETUP_RUNS, SETUP_RUNS_MIN_HALOS = 2, 1 << 62

    # -- helpers -----------------------------------------------------------------------------
    def _check_keys(self, baryonify=False):
        """HealpixRunner.py:282-287 / :407-412: models with property keys must be ParamTabulatedProfile tables."""
        keys = vars(self.model).get("p_keys", []) if self.model is not None else []
        if len(keys) > 0:
            txt = (f"You asked to use {keys} properties in Baryonification. You must pass a ParamTabulatedProfile"
                   f"as the model. You have passed {type(self.model)} instead")
            tab = getattr(self.model, "raw_input_2D", None)
            assert (not baryonify) and tab is not None and np.ndim(tab) == 3 + len(keys), txt

    def _engine(self, dev, mode, bands=None):
        """Host set-up (vectorised) + H2D + (z, M) collapse + halo binning."""
        baryonify = mode == engine.BARYONIFY
        self._check_keys(baryonify)
        t0 = time.perf_counter()
        nh = len(self.HaloLightConeCatalog.cat)
        self._progress("host set-up of %d halos" % nh)
        htab = displacement_host_table(self.model) if baryonify else profile_host_table(self.model, "2D")
        key = (mode, bands, nh)
        hint = getattr(self, "_bin_hint", {}).get(key, 0)
        extras_needed = len(vars(self.model).get("p_keys", []) or []) > 0 if self.model is not None else False
        eng = engine.ShellEngine(self.LightconeShell.NSIDE, mode, htab, None, dev, variant=self.variant, bands=bands)
        eng.max_entries = hint  # a repeated process() sizes the key list from the previous call: no counting pass ...
        if hint > 0 and self.variant == "gather" and nh >= self.SETUP_RUNS_MIN_HALOS and not extras_needed and not htab.extra:
            # ... and the catalogue goes to the device in runs of consecutive halos: upload, (z, M) collapse and binning of
            # a finished run overlap the host set-up of the next one
            eng.begin_runs(nh)
            s = host.shell_scalars(self.HaloLightConeCatalog, self.cosmo, self.mass_def, self.epsilon_max, self.model,
                                   baryonify, runs=self.SETUP_RUNS, on_run=eng.add_run)
            streamed = "_pinned" in s  # (small catalogues / no page-locked memory: nothing was sent yet)
        else:
            s = host.shell_scalars(self.HaloLightConeCatalog, self.cosmo, self.mass_def, self.epsilon_max, self.model,
                                   baryonify)
            streamed = False
        s["extras"] = host.property_columns(self.HaloLightConeCatalog.cat, self.model)
        eng.s, eng.nh = s, nh
        self.timings["host_setup"] = time.perf_counter() - t0
        after = getattr(self, "_after_host", None)
        if after is not None:  # e.g. start the map upload now: the host cores are free while the GPU works
            after()
        t0 = time.perf_counter()
        self._progress("upload, (z, M) table collapse, halo binning into sky cells")
        if streamed:
            eng.end_runs()
        else:
            eng.upload().bin()
        if eng.nentries is not None:
            self._bin_hint = {key: eng.nentries + eng.nentries // 16 + 1024}
        self._tick("upload_bin", t0, dev)
        return eng

    def _pairs(self, dev, mode):
        """Host set-up + binning + pair stage -> (engine, device accumulator); sets ``npairs``."""
        nh = len(self.HaloLightConeCatalog.cat)
        if self.variant != "gather" or self.pipeline_chunks < 2 or nh < self.pipeline_min_halos:
            eng = self._engine(dev, mode)
            t0 = time.perf_counter()
            out = eng.accumulate()
            self.npairs = eng.pair_count()
            self._tick("accumulate", t0, dev)
            return eng, out
        return self._pairs_pipelined(dev, mode, nh)

    def _pairs_pipelined(self, dev, mode, nh):
        import copy
        import queue
        import threading
        baryonify = mode == engine.BARYONIFY
        self._check_keys(baryonify)
        t00 = time.perf_counter()
        htab = displacement_host_table(self.model) if baryonify else profile_host_table(self.model, "2D")
        K = self.pipeline_chunks
        bounds = [(nh * k // K, nh * (k + 1) // K) for k in range(K)]
        q = queue.Queue()
        state = {"out": None, "eng": None, "pairs": 0, "err": None}

        def gpu_side():
            try:
                with torch.cuda.device(dev):
                    k = 0
                    while True:
                        s = q.get()
                        if s is None:
                            return
                        eng = engine.ShellEngine(self.LightconeShell.NSIDE, mode, htab, s, dev, variant="gather")
                        eng.upload().bin()
                        state["out"] = eng.accumulate(state["out"], add=k > 0)
                        state["pairs"] += eng.pair_count()  # also keeps the chunk's buffers alive until it is done
                        state["eng"] = eng
                        k += 1
            except BaseException as e:  # re-raised by the caller
                state["err"] = e

        th = threading.Thread(target=gpu_side, daemon=True)
        th.start()
        t_host = 0.0
        try:
            for i, (a, b) in enumerate(bounds):
                t0 = time.perf_counter()
                part = copy.copy(self.HaloLightConeCatalog)
                part.cat = self.HaloLightConeCatalog.cat[a:b]
                s = host.shell_scalars(part, self.cosmo, self.mass_def, self.epsilon_max, self.model, baryonify)
                s["extras"] = host.property_columns(part.cat, self.model)
                t_host += time.perf_counter() - t0
                q.put(s)
                if i == self.pipeline_upload_after:
                    after = getattr(self, "_after_host", None)
                    if after is not None:  # the map upload (host memcpy into pinned memory) joins here
                        after()
        finally:
            q.put(None)
            th.join()
        if state["err"] is not None:
            raise state["err"]
        self.npairs = state["pairs"]
        self.timings["host_setup"] = t_host
        self.timings["pairs_pipelined"] = time.perf_counter() - t00
        return state["eng"], state["out"]

    def _progress(self, what):
        """Per-stage progress lines under the reference's ``verbose`` flag (its tqdm bar over halos, HealpixRunner.py:291,
        has no per-halo counterpart here: the loop is a handful of kernel launches)."""
        if self.verbose:
            import sys
            print("[%s] %s" % (type(self).__name__, what), file=sys.stderr, flush=True)

    def _tick(self, name, t0, dev):
        torch.cuda.current_stream(dev).synchronize()  # (this stream only: the map upload on its side stream runs on)
        self.timings[name] = time.perf_counter() - t0
        return time.perf_counter()


class BaryonifyShell(DefaultRunner):
    """Baryonify a HEALPix *mass* map (RING) with a tabulated displacement model."""

    def offsets_device(self, dev):
        self._eng, off = self._pairs(dev, engine.BARYONIFY)
        return off

    PIPELINE_GROUPS = 8

    def _process_pipelined(self, dev, orig):
        """Pair stage, re-binning and read-back proceed ring-band group by group (``ShellEngine.baryonify_to_host``); the
        map travels to the device in the same pieces, starting when the host set-up has released the host cores."""
        nside = self.LightconeShell.NSIDE
        _, edges = engine.band_groups(nside, self.PIPELINE_GROUPS)
        box = {}
        # a pageable map starts its way (host memcpy into page-locked memory) as soon as the host set-up has released the
        # cores; a map that is page-locked in place has nothing to prepare and starts after the binning stage, whose small
        # uploads would otherwise stand behind it on the GPU's one host-to-device copy engine
        self._after_host = lambda: None if dv.pinned_already(orig) else box.setdefault("up", dv.ChunkedUpload(orig, dev, edges))
        try:
            eng = self._engine(dev, engine.BARYONIFY)
        finally:
            self._after_host = None
        t0 = time.perf_counter()
        up = box.get("up") or dv.ChunkedUpload(orig, dev, edges)
        self._progress("pair stage + re-binning + read-back in %d ring-band groups" % (len(edges) - 1))
        new_map, new_sum_d, old_sum_d = eng.baryonify_to_host(up, self.PIPELINE_GROUPS)
        self.npairs = eng.npairs
        self.regrid_used, self.regrid_window = eng.regrid_used, getattr(eng, "regrid_window", None)
        new_sum, old_sum = float(new_sum_d.item()), float(old_sum_d.item())
        torch.cuda.synchronize(dev)
        self.timings["pairs_regrid_d2h"] = time.perf_counter() - t0
        self.timings["pipelined"] = bool(eng.regrid_pipelined)
        return new_map, new_sum, old_sum

    def process(self):
        dev = dv.device(self.device)
        with torch.cuda.device(dev):
            orig = np.ascontiguousarray(self.LightconeShell.map, dtype=np.float64)
            nh = len(self.HaloLightConeCatalog.cat)
            if (self.variant == "gather" and self.regrid == "gather" and self.pinned_output and nh > 0
                    and not (self.pipeline_chunks > 1 and nh >= self.pipeline_min_halos)):
                new_map, new_sum, old_sum = self._process_pipelined(dev, orig)
                assert np.isclose(new_sum, old_sum), \
                    "ERROR in pixel regridding, sum(new_map) [%0.14e] != sum(oldmap) [%0.14e]" % (new_sum, old_sum)
                return new_map
            # the 8 B/pixel map travels (pageable -> pinned -> HBM) while halos are binned and evaluated on the GPU; it
            # starts after the host set-up, which wants all host cores for itself
            box = {}
            self._after_host = lambda: None if dv.pinned_already(orig) else box.setdefault("up", dv.AsyncUpload(orig, dev))
            try:
                off = self.offsets_device(dev)
            finally:
                self._after_host = None
            t0 = time.perf_counter()
            orig_d = (box.get("up") or dv.AsyncUpload(orig, dev)).result()
            old_sum_d = orig_d.sum()  # the reference's two np.sum's (HealpixRunner.py:344-346), on the device
            # re-binning and read-back overlap: the gather variant hands over the map range by range
            self._progress("re-binning (%s)" % self.regrid)
            new_map, new_sum_d, t_dev = self._eng.regrid_to_host(orig_d, off, method=self.regrid, pinned=self.pinned_output)
            self.regrid_used = self._eng.regrid_used
            self._eng = None
            new_sum, old_sum = float(new_sum_d.item()), float(old_sum_d.item())
            torch.cuda.synchronize(dev)
            t_all = time.perf_counter() - t0
            self.timings["regrid"], self.timings["d2h"] = t_dev, max(t_all - t_dev, 0.0)  # (d2h: what the copy adds)
        assert np.isclose(new_sum, old_sum), \
            "ERROR in pixel regridding, sum(new_map) [%0.14e] != sum(oldmap) [%0.14e]" % (new_sum, old_sum)
        return new_map


class PaintProfilesShell(DefaultRunner):
    """Paint a tabulated projected profile (e.g. tSZ) around every halo onto a HEALPix map (RING)."""

    def process_device(self, dev):
        assert self.model is not None, "You must provide a model"
        _, out = self._pairs(dev, engine.PAINT)
        return out

    def process(self):
        dev = dv.device(self.device)
        with torch.cuda.device(dev):
            nh = len(self.HaloLightConeCatalog.cat)
            if (not self.pinned_output or self.variant != "gather"
                    or (self.pipeline_chunks > 1 and nh >= self.pipeline_min_halos)):
                return dv.to_host(self.process_device(dev), self.pinned_output)
            # pixel owners finish band by band: the read-back of a finished band group overlaps the next group's kernel
            assert self.model is not None, "You must provide a model"
            eng = self._engine(dev, engine.PAINT)
            t0 = time.perf_counter()
            out = eng.accumulate_to_host()
            self.npairs = eng.npairs
            self._tick("accumulate_d2h", t0, dev)
            return out
