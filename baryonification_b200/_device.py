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

    def collapse(self, z_in, M_in):
        """rows[j, :] for every halo (bf_table_collapse_zM)."""
        nh = z_in.numel()
        rows = empty((nh, self.nr), dev=z_in.device)
        _lib.call("bf_table_collapse_zM", C.byref(self.desc), nh, ptr(z_in), ptr(M_in), ptr(rows), stream_ptr())
        return rows

    def readout(self, row, r, lnr_shift=0.0, cut=0.0, mode=0):
        out = torch.empty_like(r)
        _lib.call("bf_table_readout", C.byref(self.desc), ptr(row), r.numel(), ptr(r), float(lnr_shift), float(cut),
                  int(mode), ptr(out), stream_ptr())
        return out


def pair_counter(dev=None):
    return torch.zeros(1, dtype=torch.int64, device=dev or device())
