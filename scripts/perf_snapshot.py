"""BaryonifySnapshot stage timings.  Usage: perf_snapshot.py [NPART_SIDE] [NHALO] [DIM]"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import ref_shim

ref_shim.install()
import baryonification_b200 as bf
from baryonification_b200 import _device as dv, _lib, synthetic as syn

side = int(sys.argv[1]) if len(sys.argv) > 1 else 512
nh = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
dim = int(sys.argv[3]) if len(sys.argv) > 3 else 3
t0 = time.perf_counter()
cfg = syn.config_snapshot(npart_side=side, nhalo=nh, dim=dim)
print(f"inputs {time.perf_counter()-t0:.1f}s", flush=True)
r = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], cfg["epsilon_max"], cfg["model"], verbose=False, device=0)
dev = dv.device(0)
cat = cfg["field"].cat
npart = cat.size
for i in range(3):
    t0 = time.perf_counter()
    out = r.process()
    torch.cuda.synchronize()
    print(f"process() {1e3*(time.perf_counter()-t0):9.1f} ms  pairs {r.npairs:.4e}  entries {r.nentries}  "
          f"stages {({k: round(v*1e3,1) for k,v in r.timings.items()})}", flush=True)
# device-resident stages
def ev(fn, n=3):
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts)
t_setup = ev(lambda: r._setup(dev))
st = r._setup(dev)
rec = dv.to_dev(cat.view(np.float64).reshape(-1, 4), dev=dev)
rec_out = torch.empty_like(rec)
cnt = dv.pair_counter(dev)
def k_rec():
    _lib.call("bf_snap_displace_records", st["dim"], npart, dv.ptr(rec), st["L"], float(st["s"]["a"]), C.byref(st["desc"]),
              C.byref(st["table"].desc), st["ng"], dv.ptr(st["ws"]), dv.ptr(rec_out), dv.ptr(cnt), dv.stream_ptr())
t_rec = ev(k_rec)
xs = [rec[:, i].contiguous() for i in (1, 2, 3)]
xo = [torch.empty_like(v) for v in xs]
def k_soa():
    _lib.call("bf_snap_displace", st["dim"], npart, dv.ptr(xs[0]), dv.ptr(xs[1]), dv.ptr(xs[2] if st["dim"] == 3 else None),
              st["L"], float(st["s"]["a"]), C.byref(st["desc"]), C.byref(st["table"].desc), st["ng"], dv.ptr(st["ws"]),
              dv.ptr(xo[0]), dv.ptr(xo[1]), dv.ptr(xo[2] if st["dim"] == 3 else None), dv.ptr(cnt), dv.stream_ptr())
t_soa = ev(k_soa)
print(f"ng {st['ng']}  setup+bin (host incl.) {t_setup:.2f} ms   records kernel {t_rec:.3f} ms "
      f"({64*npart/t_rec*1e-6:.0f} GB/s of 64 B/particle; {88*npart/t_rec*1e-6:.0f} GB/s at the reference's 88 B/particle)   "
      f"SoA kernel {t_soa:.3f} ms ({(16*st['dim'])*npart/t_soa*1e-6:.0f} GB/s)", flush=True)
