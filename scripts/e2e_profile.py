"""Fine-grained wall-clock breakdown of BaryonifyShell.process() (host buffers in, host map out).
Usage: e2e_profile.py [NSIDE] [NHALO]"""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import ref_shim

ref_shim.install()
import baryonification_b200 as bf
from baryonification_b200 import synthetic as syn

nside = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nh = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
cfg = syn.config_shell(nside=nside, nhalo=nh)
r = bf.BaryonifyShell(cfg["catalog"], cfg["field"], cfg["epsilon_max"], cfg["model"], verbose=False, device=0)
for i in range(4):
    t0 = time.perf_counter()
    out = r.process()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"call {i}: {dt*1e3:8.1f} ms  stages", {k: round(v * 1e3, 1) for k, v in r.timings.items()},
          "sum", round(sum(r.timings.values()) * 1e3, 1), flush=True)
pr = cProfile.Profile()
pr.enable()
out = r.process()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
