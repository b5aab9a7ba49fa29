"""Device-side stage times of BaryonifyGrid (CUDA events): pair kernel and the three re-binning variants.
Usage: perf_grid.py [dim] [npix] [nhalo]"""
import ctypes as C
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle import ref_shim  # noqa: E402  (inputs only: pyccl stand-in)
ref_shim.install()
import baryonification_b200 as bf  # noqa: E402
from baryonification_b200 import _device as dv, _lib, synthetic as syn  # noqa: E402

dim = int(sys.argv[1]) if len(sys.argv) > 1 else 3
npix = int(sys.argv[2]) if len(sys.argv) > 2 else 512
nhalo = int(sys.argv[3]) if len(sys.argv) > 3 else 20000
cfg = syn.config_grid(npix=npix, nhalo=nhalo, dim=dim, epsilon_max=20.0)
r = bf.BaryonifyGrid(cfg["catalog"], cfg["field"], 20.0, cfg["model"], verbose=False)
dev = dv.device(None)
lib = _lib.load(require_cuda=True)


def ev(fn, n=5):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts)


for variant in ("atomic", "gather"):
    r.variant = variant
    for _ in range(2):
        off = r.offsets_device(dev)
    print(variant, "pairs", r.npairs, "accumulate (host clock: bin + pair kernel) ms %.2f" % (r.timings["accumulate"] * 1e3),
          "entries", getattr(r, "nentries", None))
orig = dv.to_dev(np.ascontiguousarray(cfg["field"].map).reshape(-1))
new = dv.zeros(orig.numel())
ncell = orig.numel()
ws_g = torch.empty(64, dtype=torch.uint8, device=dev)
win = (C.c_int * 3)()
t = ev(lambda: _lib.check(lib.bf_grid_regrid_gather(dim, npix, dv.ptr(off), dv.ptr(orig), dv.ptr(new), dv.ptr(ws_g), 64, win,
                                                    dv.stream_ptr()), "gather"))
print("regrid gather ms %.3f  window %s  -> %.1f GB/s on %d B/cell" % (t, tuple(win), ncell * (64 if dim == 2 else 168) / t / 1e6,
                                                                    64 if dim == 2 else 168))
t = ev(lambda: _lib.call("bf_grid_regrid", dim, npix, dv.ptr(off), dv.ptr(orig), dv.ptr(new), dv.stream_ptr()))
print("regrid atomic ms %.3f" % t)
if (ncell << dim) < 2**31:
    nb = lib.bf_grid_regrid_sorted_ws_bytes(dim, npix)
    ws = torch.empty(nb, dtype=torch.uint8, device=dev)
    t = ev(lambda: _lib.call("bf_grid_regrid_sorted", dim, npix, dv.ptr(off), dv.ptr(orig), dv.ptr(new), dv.ptr(ws), nb,
                             dv.stream_ptr()), 3)
    print("regrid sorted ms %.3f (workspace %.1f GB)" % (t, nb / 1e9))
for k in range(3):
    t0 = time.perf_counter(); out = r.process(); t1 = time.perf_counter()
    print("process() ms %.1f" % ((t1 - t0) * 1e3), {k: round(v * 1e3, 1) for k, v in r.timings.items()}, r.regrid_used)
