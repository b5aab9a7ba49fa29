"""Delta P(k): the power-spectrum suppression of a baryonified particle set on the GPU.

Mirrors the notebook recipe of the reference (``examples/10_Reproduce_Schneider_deltaPk.ipynb``): cell 12 sets
``kbins = linspace(k_f, k_f * Ngrd / 2, Nk + 1)`` and ``kinds = floor((k - kbins[0]) / (kbins[1] - kbins[0]))``; cells
15-21 fold the positions (``Part % (L / factor)``), histogram them on ``Ngrd^3`` cells (cell 1), ``np.fft.fftn``,
``|F|^2`` and ``np.bincount`` per bin, and report ``S(k) = P_baryon / P_dmo`` at ``k_cen * factor``.
The FFT is cuFFT through ``torch.fft.rfftn`` (a library transform, as BASELINE.json's config 5 prescribes); the
deposit and the deterministic k-binning are this package's kernels (csrc/bf_pk.cu).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _device as dv
from . import _lib


class PowerSpectrumGrid:
    def __init__(self, Ngrd, Lbox, Nk=180, device=None):
        self.Ngrd, self.Lbox, self.Nk, self.device = int(Ngrd), float(Lbox), int(Nk), device
        self.npairs = None

    def _setup(self, factor):
        """(fold length, fftfreq scale, kbins[0], kbins[1] - kbins[0]).  The k bins are those of notebook cell 12, which
        computes them ONCE for the whole box (``Lbox = Snap.L``); the ``for factor`` loop of cells 15-21 only folds the
        positions into ``Snap.L / factor`` and reports ``k_cen * factor`` -- so the binning constants do not depend on
        ``factor`` (pinned against the notebook's own code: tests/golden/make_golden_pk.py)."""
        L = self.Lbox
        kbins = np.linspace(2 * np.pi / L, 2 * np.pi / L * self.Ngrd / 2, self.Nk + 1)
        d = 1 / (2 * np.pi / L) / self.Ngrd
        val = 1.0 / (self.Ngrd * d)  # np.fft.fftfreq(n, d): results * val
        return self.Lbox / factor, float(val), float(kbins[0]), float(kbins[1] - kbins[0])

    def upload(self, particles, dev=None):
        """Host particles -> a device form ``density`` accepts, so that several folding factors read one upload:
        the (M, x, y, z) record tensor when the array has the reference's 32-byte layout, else (x, y, z) columns."""
        from .Runners.snapshot import BaryonifySnapshot
        dev = dev or dv.device(self.device)
        if isinstance(particles, (torch.Tensor, tuple)):
            return particles
        if getattr(particles, "dtype", None) is not None and particles.dtype.names:
            if BaryonifySnapshot._is_reference_layout(particles):
                return dv.AsyncUpload(particles.view(np.float64).reshape(-1, 4), dev).result()
            cols = [particles[k] for k in "xyz"]
        else:
            p = np.asarray(particles)
            cols = [p[:, 0], p[:, 1], p[:, 2]]
        return tuple(dv.to_dev(c, dev=dev, pinned=True) for c in cols)

    def density(self, particles, factor=1, dev=None):
        """NGP counts [Ngrd]^3 (float64, device) of ``particles`` folded into a box of L / factor.
        ``particles``: structured array with x, y, z (the runners' output), an [n, 3] array, or a device tensor
        of (M, x, y, z) records [n, 4]."""
        dev = dev or dv.device(self.device)
        L, _, _, _ = self._setup(factor)
        grid = dv.zeros((self.Ngrd,) * 3, dev=dev)
        with torch.cuda.device(dev):
            particles = self.upload(particles, dev)
            if isinstance(particles, torch.Tensor):
                rec = particles.contiguous()
                assert rec.dim() == 2 and rec.shape[1] == 4 and rec.dtype == torch.float64
                _lib.call("bf_pk_ngp_deposit", rec.shape[0], None, None, None, dv.ptr(rec), L, self.Ngrd, dv.ptr(grid),
                          dv.stream_ptr())
            else:
                x, y, z = particles
                _lib.call("bf_pk_ngp_deposit", x.numel(), dv.ptr(x), dv.ptr(y), dv.ptr(z), None, L, self.Ngrd,
                          dv.ptr(grid), dv.stream_ptr())
        return grid

    def power(self, grid, factor=1):
        """(k_cen * factor, P(k), modes per bin) of a density grid: sum |F|^2 / N_modes in each of the Nk bins."""
        dev = grid.device
        _, val, k0, dk = self._setup(factor)
        with torch.cuda.device(dev):
            F = torch.fft.rfftn(grid)  # cuFFT
            lib = _lib.load(require_cuda=True)
            nb = lib.bf_pk_bin_ws_bytes(self.Ngrd, self.Nk)
            ws = torch.empty(nb, dtype=torch.uint8, device=dev)
            out = dv.empty((3, self.Nk), dev=dev)
            strides = (C.c_int64 * 3)(*F.stride())  # torch.fft hands back a permuted dense view: read it in place
            _lib.call("bf_pk_bin", self.Ngrd, dv.ptr(F), strides, val, k0, dk, self.Nk, dv.ptr(out), dv.ptr(ws), nb,
                      dv.stream_ptr())
            o = out.cpu().numpy()
        with np.errstate(divide="ignore", invalid="ignore"):
            return o[1] / o[2] * factor, o[0] / o[2], o[2]

    def suppression(self, part_baryon, part_dmo, factors=(1, 8)):
        """{factor: {'k': k_cen * factor, 'S': P_baryon / P_dmo}} as the notebook's RES_* dictionaries."""
        res = {}
        part_baryon, part_dmo = self.upload(part_baryon), self.upload(part_dmo)  # once, for every factor
        for f in factors:
            k, PB, _ = self.power(self.density(part_baryon, f), f)
            _, PD, _ = self.power(self.density(part_dmo, f), f)
            res[f"Factor{f}"] = {"k": k, "S": PB / PD}
        return res
