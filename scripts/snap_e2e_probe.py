"""BaryonifySnapshot.process() end to end at the bench size, stage timings, staged vs page-locked-in-place input."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import baryonification_b200 as bf
from baryonification_b200 import synthetic as syn
cfg = syn.config_snapshot(npart_side=512, nhalo=100_000, dim=3)
r = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], cfg["epsilon_max"], cfg["model"], verbose=False, device=0)
for i in range(6):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = r.process()
    dt = time.perf_counter() - t0
    print(os.environ.get("BF_PIN_INPUTS", "default"), "call", i, "%.1f ms" % (dt * 1e3), {k: round(v * 1e3, 1) for k, v in r.timings.items()}, flush=True)
