"""Host-side copy rates on the GPU box: pageable -> page-locked staging with 1..16 torch threads (main and helper
thread), numpy copies in a thread pool, H2D / D2H from torch-pinned memory and from a cudaHostRegister'ed /dev/shm map."""
import os, sys, time, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from concurrent.futures import ThreadPoolExecutor

n = 100_000_000  # 800 MB
src = torch.from_numpy(np.random.default_rng(0).random(n))
pin = torch.empty(n, dtype=torch.float64, pin_memory=True)
print("cpus", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)), "torch threads", torch.get_num_threads(), flush=True)

def rate(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); best = min(best, time.perf_counter() - t0)
    return 0.8 / best

for nt in (1, 2, 4, 8, 16):
    def work():
        torch.set_num_threads(nt)
        print("helper thread torch copy_ %2d threads: %.1f GB/s" % (nt, rate(lambda: pin.copy_(src))), flush=True)
    th = threading.Thread(target=work); th.start(); th.join()
for nt in (1, 2, 4, 8, 16):
    pool = ThreadPoolExecutor(nt)
    d, s = pin.numpy(), src.numpy()
    edges = np.linspace(0, n, nt + 1).astype(np.int64)
    def go():
        list(pool.map(lambda i: np.copyto(d[edges[i]:edges[i + 1]], s[edges[i]:edges[i + 1]]), range(nt)))
    print("numpy copyto in %2d pool threads: %.1f GB/s" % (nt, rate(go)), flush=True)
if torch.cuda.is_available():
    dev = torch.device("cuda", 0)
    d_t = torch.empty(n, dtype=torch.float64, device=dev)
    def h2d(): d_t.copy_(pin, non_blocking=True); torch.cuda.synchronize()
    def d2h(): pin.copy_(d_t, non_blocking=True); torch.cuda.synchronize()
    print("H2D pinned %.1f GB/s, D2H pinned %.1f GB/s" % (rate(h2d), rate(d2h)), flush=True)
    from baryonification_b200.sharding import SharedHostArena
    ar = SharedHostArena(n)
    t = ar.tensor
    def d2h_shm(): t.copy_(d_t, non_blocking=ar.registered); torch.cuda.synchronize()
    def h2d_shm(): d_t.copy_(t, non_blocking=ar.registered); torch.cuda.synchronize()
    print("registered /dev/shm arena (registered=%s): H2D %.1f GB/s, D2H %.1f GB/s" % (ar.registered, rate(h2d_shm), rate(d2h_shm)), flush=True)
    def pageable_h2d(): d_t.copy_(src); torch.cuda.synchronize()
    print("H2D straight from pageable memory: %.1f GB/s" % rate(pageable_h2d), flush=True)
