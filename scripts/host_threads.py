"""Wall time of host.shell_scalars vs BF_HOST_THREADS (run on the GPU box's host)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim
ref_shim.install()
from baryonification_b200 import synthetic as syn, host
from baryonification_b200._ccl import default_mass_def
cfg = syn.config_shell(nside=64, nhalo=1_000_000)
for nt in (1, 2, 4, 8, 16, 32):
    os.environ["BF_HOST_THREADS"] = str(nt)
    host._pool = None
    ts = []
    for _ in range(4):
        t = time.perf_counter()
        host.shell_scalars(cfg["catalog"], cfg["catalog"].cosmo, default_mass_def(), 20.0, cfg["model"], True)
        ts.append(time.perf_counter() - t)
    print(nt, [round(x * 1e3, 1) for x in ts], flush=True)
