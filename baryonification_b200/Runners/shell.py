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
from .. import _lib, host
from .._ccl import default_mass_def
from ..tables import displacement_host_table, profile_host_table

__all__ = ["DefaultRunner", "BaryonifyShell", "PaintProfilesShell"]


class DefaultRunner:
    def __init__(self, HaloLightConeCatalog, LightconeShell, epsilon_max, model, use_ellipticity=False,
                 mass_def=None, verbose=True, *, device=None, variant="atomic"):
        self.HaloLightConeCatalog = HaloLightConeCatalog
        self.LightconeShell = LightconeShell
        self.cosmo = HaloLightConeCatalog.cosmology
        self.model = model
        self.epsilon_max = epsilon_max
        self.mass_def = mass_def if mass_def is not None else default_mass_def()
        self.verbose = verbose
        self.use_ellipticity = use_ellipticity
        self.device = device
        self.variant = variant
        self.npairs = None
        self.timings = {}
        if use_ellipticity:  # HealpixRunner.py:164-165
            raise NotImplementedError("You have set use_ellipticity = True, but this not yet implemented for HealpixRunner")

    # -- helpers -----------------------------------------------------------------------------
    def _check_keys(self):
        keys = vars(self.model).get("p_keys", []) if self.model is not None else []
        if len(keys) > 0:  # HealpixRunner.py:282-287; property-keyed tables are not on the GPU path yet
            raise NotImplementedError(f"You asked to use {keys} properties in Baryonification; "
                                      "ParamTabulatedProfile tables are not yet supported by the B200 runners")

    def _upload_halos(self, s, table, dev, baryonify):
        """H2D of the per-halo arrays + the (z, M) collapse; returns (ShellHalos descriptor, keep-alive list)."""
        t = {k: dv.to_dev(s[k], dev=dev) for k in ("vec", "lon", "lat", "radius", "D", "a", "cut", "lnr_shift",
                                                     "z_in", "M_in")}
        order = torch.argsort(t["radius"], descending=True)  # largest discs first (load balance)
        rows = table.collapse(t["z_in"], t["M_in"])
        desc = _lib.ShellHalos(t["vec"].data_ptr(), t["lon"].data_ptr(), t["lat"].data_ptr(), t["radius"].data_ptr(),
                               t["D"].data_ptr(), t["a"].data_ptr(), t["cut"].data_ptr(), t["lnr_shift"].data_ptr(),
                               rows.data_ptr(), order.data_ptr(), s["a"].size)
        return desc, (t, order, rows)

    def _tick(self, name, t0, dev):
        torch.cuda.synchronize(dev)
        self.timings[name] = time.perf_counter() - t0
        return time.perf_counter()


class BaryonifyShell(DefaultRunner):
    """Baryonify a HEALPix *mass* map (RING) with a tabulated displacement model."""

    def offsets_device(self, dev):
        self._check_keys()
        t0 = time.perf_counter()
        table = dv.DeviceTable(displacement_host_table(self.model), dev=dev)
        s = host.shell_scalars(self.HaloLightConeCatalog, self.cosmo, self.mass_def, self.epsilon_max, self.model, True)
        t0 = self._tick("host_setup", t0, dev)
        NSIDE = self.LightconeShell.NSIDE
        desc, keep = self._upload_halos(s, table, dev, True)
        off = dv.zeros((12 * NSIDE * NSIDE, 3), dev=dev)
        cnt = dv.pair_counter(dev)
        _lib.call("bf_shell_baryonify_atomic", NSIDE, C.byref(desc), C.byref(table.desc), dv.ptr(off), dv.ptr(cnt),
                  dv.stream_ptr())
        self._tick("accumulate", t0, dev)
        self.npairs = int(cnt.item())
        del keep
        return off

    def process(self):
        dev = dv.device(self.device)
        with torch.cuda.device(dev):
            orig = np.ascontiguousarray(self.LightconeShell.map, dtype=np.float64)
            NSIDE = self.LightconeShell.NSIDE
            off = self.offsets_device(dev)
            t0 = time.perf_counter()
            orig_d = dv.to_dev(orig, dev=dev, pinned=True)
            new_d = dv.zeros(orig.size, dev=dev)
            _lib.call("bf_shell_regrid", NSIDE, dv.ptr(orig_d), dv.ptr(off), dv.ptr(new_d), dv.stream_ptr())
            new_sum = float(new_d.sum().item())
            new_map = new_d.cpu().numpy()
            self._tick("regrid", t0, dev)
        old_sum = np.sum(orig)  # HealpixRunner.py:344-346
        assert np.isclose(new_sum, old_sum), \
            "ERROR in pixel regridding, sum(new_map) [%0.14e] != sum(oldmap) [%0.14e]" % (new_sum, old_sum)
        return new_map


class PaintProfilesShell(DefaultRunner):
    """Paint a tabulated projected profile (e.g. tSZ) around every halo onto a HEALPix map (RING)."""

    def process_device(self, dev):
        self._check_keys()
        assert self.model is not None, "You must provide a model"
        t0 = time.perf_counter()
        table = dv.DeviceTable(profile_host_table(self.model, "2D"), dev=dev)
        s = host.shell_scalars(self.HaloLightConeCatalog, self.cosmo, self.mass_def, self.epsilon_max, self.model, False)
        t0 = self._tick("host_setup", t0, dev)
        NSIDE = self.LightconeShell.NSIDE
        desc, keep = self._upload_halos(s, table, dev, False)
        out = dv.zeros(12 * NSIDE * NSIDE, dev=dev)
        cnt = dv.pair_counter(dev)
        _lib.call("bf_shell_paint_atomic", NSIDE, C.byref(desc), C.byref(table.desc), dv.ptr(out), dv.ptr(cnt),
                  dv.stream_ptr())
        self._tick("accumulate", t0, dev)
        self.npairs = int(cnt.item())
        del keep
        return out

    def process(self):
        dev = dv.device(self.device)
        with torch.cuda.device(dev):
            return self.process_device(dev).cpu().numpy()
