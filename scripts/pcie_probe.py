"""Host-side transfer rates on the GPU box: pageable->pinned memcpy (torch copy_, numpy), H2D, D2H, both at once."""
import time, numpy as np, torch
n = 1 << 27  # 1 GiB of f64
src = np.random.default_rng(0).random(n)
pin = torch.empty(n, dtype=torch.float64, pin_memory=True)
pin2 = torch.empty(n, dtype=torch.float64, pin_memory=True)
d = torch.empty(n, dtype=torch.float64, device="cuda")
d2 = torch.empty(n, dtype=torch.float64, device="cuda")
t = torch.from_numpy(src)
def tm(f, k=3):
    f(); torch.cuda.synchronize(); best = 1e9
    for _ in range(k):
        t0 = time.perf_counter(); f(); torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
    return n * 8 / best / 1e9
print("threads", torch.get_num_threads())
print("torch copy_ pageable->pinned GB/s", tm(lambda: pin.copy_(t)))
print("numpy copyto pageable->pinned GB/s", tm(lambda: np.copyto(pin.numpy(), src)))
print("H2D pinned GB/s", tm(lambda: d.copy_(pin, non_blocking=True)))
print("D2H pinned GB/s", tm(lambda: pin2.copy_(d2, non_blocking=True)))
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def both():
    with torch.cuda.stream(s1): d.copy_(pin, non_blocking=True)
    with torch.cuda.stream(s2): pin2.copy_(d2, non_blocking=True)
print("H2D+D2H concurrent, GB/s each way", tm(both))
import threading
def par(k):
    def f():
        th = [threading.Thread(target=lambda i=i: pin[i * n // k:(i + 1) * n // k].copy_(t[i * n // k:(i + 1) * n // k])) for i in range(k)]
        [x.start() for x in th]; [x.join() for x in th]
    return f
for k in (2, 4):
    print("torch copy_ in %d python threads GB/s" % k, tm(par(k)))
print("H2D from pageable GB/s", tm(lambda: d.copy_(t)))
