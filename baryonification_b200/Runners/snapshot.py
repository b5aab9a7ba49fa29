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
    """``process()`` returns a displaced copy of ``ParticleSnapshot.cat`` (SnapshotRunner.py:250-262)."""

    @staticmethod
    def _is_reference_layout(cat):
        """(M, x, y, z) native float64 records of 32 bytes (utils/io.py:581-589): the kernel reads them as they are."""
        d = cat.dtype
        if d.names is None or d.itemsize != 32 or not cat.flags.c_contiguous:
            return False
        want = {"M": 0, "x": 8, "y": 16, "z": 24}
        return all(k in d.fields and d.fields[k][1] == o and d.fields[k][0] == np.dtype("=f8") for k, o in want.items())

    def _setup(self, dev):
        """Table, per-halo arrays, (z, M) collapse and halo binning into the periodic grid."""
        S = self.ParticleSnapshot
        if len(vars(self.model).get("p_keys", [])) > 0:
            raise NotImplementedError("property-keyed tables are not yet supported by the B200 runners")
        lib = _lib.load(require_cuda=True)
        table = dv.DeviceTable(displacement_host_table(self.model), dev=dev)
        s = host.snapshot_scalars(self.HaloNDCatalog, S, self.mass_def, self.epsilon_max, self.model)
        dim, nh = s["dim"], s["Rq"].size
        t = {k: dv.to_dev(s[k], dev=dev) for k in ("pos", "Rq", "cut", "lnr_shift", "z_in", "M_in")}
        rows = table.collapse(t["z_in"], t["M_in"])
        desc = _lib.SnapHalos(t["pos"].data_ptr(), t["Rq"].data_ptr(), t["cut"].data_ptr(),
                              t["lnr_shift"].data_ptr(), rows.data_ptr(), nh)
        L = float(S.L)
        ng = lib.bf_snap_grid_cells(dim, L, float(s["Rq"].mean()) if nh else 0.0)
        if os.environ.get("BF_SNAP_NG"):  # tuning experiments only
            ng = int(os.environ["BF_SNAP_NG"])
        max_entries, ws = 0, None
        for _ in range(2):  # counting pass first (exact size), then the real one
            need = lib.bf_snap_ws_bytes(dim, nh, L, ng, max_entries)
            ws = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
            n = C.c_int64()
            rc = lib.bf_snap_bin_halos(dim, L, ng, C.byref(desc), max_entries, dv.ptr(ws), ws.numel(), C.byref(n),
                                       dv.stream_ptr())
            if rc == -2 and n.value > max_entries:
                max_entries = n.value
                continue
            _lib.check(rc, "bf_snap_bin_halos")
            break
        else:
            raise _lib.BFError("bf_snap_bin_halos: workspace retry failed")
        self.nentries = n.value
        return dict(table=table, s=s, t=t, rows=rows, desc=desc, ng=ng, ws=ws, dim=dim, L=L)

    def process(self):
        dev = dv.device(self.device)
        S = self.ParticleSnapshot
        with torch.cuda.device(dev):
            t0 = time.perf_counter()
            cat = S.cat
            fast = self._is_reference_layout(cat)
            up = dv.AsyncUpload(cat.view(np.float64).reshape(-1, 4), dev) if fast else None  # travels during set-up
            st = self._setup(dev)
            dim, s = st["dim"], st["s"]
            cnt = dv.pair_counter(dev)
            self.timings["setup"] = time.perf_counter() - t0
            t1 = time.perf_counter()
            if fast:
                rec = up.result()
                rec_out = torch.empty_like(rec)
                _lib.call("bf_snap_displace_records", dim, rec.shape[0], dv.ptr(rec), st["L"], float(s["a"]),
                          C.byref(st["desc"]), C.byref(st["table"].desc), st["ng"], dv.ptr(st["ws"]), dv.ptr(rec_out),
                          dv.ptr(cnt), dv.stream_ptr())
                new_cat = dv.to_host(rec_out, True).reshape(-1).view(cat.dtype).reshape(cat.shape)
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
