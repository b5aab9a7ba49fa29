"""BaryonifyShell.process() wall time at the bench workload for several halo-run pipeline settings."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import ref_shim
ref_shim.install()
import baryonification_b200 as bf
from baryonification_b200 import synthetic as syn
cfg = syn.config_shell(nside=4096, nhalo=1_000_000)
for K, after in ((1, 0), (2, 1), (2, 0), (3, 2), (4, 3), (4, 1), (6, 5), (8, 7)):
    r = bf.BaryonifyShell(cfg["catalog"], cfg["field"], 20.0, cfg["model"], verbose=False)
    r.pipeline_chunks, r.pipeline_upload_after = K, after
    keep = [r.process() for _ in range(2)]
    ts = []
    for _ in range(4):
        torch.cuda.synchronize(); t0 = time.perf_counter(); out = r.process(); ts.append(time.perf_counter() - t0)
    print(K, after, "ms", [round(t * 1e3, 1) for t in ts], {k: round(v * 1e3, 1) for k, v in r.timings.items()}, flush=True)
    del keep, out
