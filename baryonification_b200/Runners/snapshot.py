"""Particle-snapshot runner on the GPU: ``BaryonifySnapshot``.

Drop-in for ``BaryonForge/Runners/SnapshotRunner.py`` (``DefaultRunnerSnapshot`` :21-155,
``BaryonifySnapshot`` :159-262).  ``KDTree_kwargs`` is accepted and ignored: the periodic ball query
is a cell-list gather on the device, there is no tree to build.
"""
from __future__ import annotations

import ctypes as C
import time

import numpy as np
import torch

from .. import _device as dv
from .. import _lib, host
from .._ccl import default_mass_def
from ..tables import displacement_host_table

__all__ = ["DefaultRunnerSnapshot", "BaryonifySnapshot"]


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


class BaryonifySnapshot(DefaultRunnerSnapshot):
    def process(self):
        dev = dv.device(self.device)
        S = self.ParticleSnapshot
        if len(vars(self.model).get("p_keys", [])) > 0:
            raise NotImplementedError("property-keyed tables are not yet supported by the B200 runners")
        with torch.cuda.device(dev):
            t0 = time.perf_counter()
            table = dv.DeviceTable(displacement_host_table(self.model), dev=dev)
            s = host.snapshot_scalars(self.HaloNDCatalog, S, self.mass_def, self.epsilon_max, self.model)
            dim, nh = s["dim"], s["Rq"].size
            t = {k: dv.to_dev(s[k], dev=dev) for k in ("pos", "Rq", "cut", "lnr_shift", "z_in", "M_in")}
            rows = table.collapse(t["z_in"], t["M_in"])
            desc = _lib.SnapHalos(t["pos"].data_ptr(), t["Rq"].data_ptr(), t["cut"].data_ptr(),
                                  t["lnr_shift"].data_ptr(), rows.data_ptr(), nh)
            names = ("x", "y", "z")[:dim]
            xin = [dv.to_dev(S.cat[k], dev=dev, pinned=True) for k in names]
            xout = [torch.empty_like(v) for v in xin]
            Rq_max = float(s["Rq"].max()) if nh else 0.0
            lib = _lib.load(require_cuda=True)
            ws_bytes = lib.bf_snap_ws_bytes(dim, nh, float(S.L), Rq_max)
            ws = torch.empty(max(ws_bytes, 256), dtype=torch.uint8, device=dev)
            cnt = dv.pair_counter(dev)
            _lib.call("bf_snap_displace", dim, xin[0].numel(), dv.ptr(xin[0]), dv.ptr(xin[1]),
                      dv.ptr(xin[2] if dim == 3 else None), float(S.L), float(s["a"]), C.byref(desc),
                      C.byref(table.desc), Rq_max, dv.ptr(xout[0]), dv.ptr(xout[1]),
                      dv.ptr(xout[2] if dim == 3 else None), dv.ptr(cnt), dv.ptr(ws), ws_bytes, dv.stream_ptr())
            self.npairs = int(cnt.item())
            new_cat = S.cat.copy()  # SnapshotRunner.py:250
            for k, v in zip(names, xout):
                new_cat[k] = v.cpu().numpy()
            self.timings["process"] = time.perf_counter() - t0
        return new_cat
