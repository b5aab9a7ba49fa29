"""H2D from a cudaHostRegister'ed ordinary numpy array (anonymous 4 KB pages) vs torch page-locked memory, alone and with a
concurrent D2H (full duplex)."""
import os, sys, time, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from baryonification_b200 import _lib
lib = _lib.load(require_cuda=True)
n = 128 * 1024 * 1024  # 1 GiB of float64
a = np.random.default_rng(0).random(n)
dev = torch.device("cuda", 0)
d_in, d_out = torch.empty(n, dtype=torch.float64, device=dev), torch.ones(n, dtype=torch.float64, device=dev)
pin_in, pin_out = torch.empty(n, dtype=torch.float64, pin_memory=True), torch.empty(n, dtype=torch.float64, pin_memory=True)
pin_in.copy_(torch.from_numpy(a))
t0 = time.perf_counter(); rc = lib.bf_host_pin(C.c_void_p(a.ctypes.data), a.nbytes); print("bf_host_pin rc", rc, "%.1f ms for 1 GiB" % ((time.perf_counter() - t0) * 1e3))
reg = torch.from_numpy(a)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def rate(fn, nbytes, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
    return nbytes / best / 1e9
GB = n * 8
print("H2D torch-pinned        %.1f GB/s" % rate(lambda: d_in.copy_(pin_in, non_blocking=True), GB))
print("H2D registered numpy    %.1f GB/s" % rate(lambda: d_in.copy_(reg, non_blocking=True), GB))
def duplex(src):
    with torch.cuda.stream(s1): d_in.copy_(src, non_blocking=True)
    with torch.cuda.stream(s2): pin_out.copy_(d_out, non_blocking=True)
print("duplex torch-pinned in + pinned out   %.1f GB/s total" % rate(lambda: duplex(pin_in), 2 * GB))
print("duplex registered in + pinned out     %.1f GB/s total" % rate(lambda: duplex(reg), 2 * GB))
print("D2H pinned alone        %.1f GB/s" % rate(lambda: pin_out.copy_(d_out, non_blocking=True), GB))
lib.bf_host_unpin(C.c_void_p(a.ctypes.data))
