"""Where the host set-up of BaryonifyShell goes (run on the GPU box's host): fixed part vs per-halo part, threads."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from baryonification_b200 import synthetic as syn, host
from baryonification_b200._ccl import default_mass_def, get_ccl
from scipy import interpolate
cfg = syn.config_shell(nside=64, nhalo=1_000_000)
cat, model, md = cfg["catalog"], cfg["model"], default_mass_def()
def best(fn, n=5):
    ts = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    return min(ts) * 1e3
print("cpus", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)))
ccl = get_ccl()
print("runner_cosmology %.2f ms" % best(lambda: host.runner_cosmology(cat.cosmology, True)))
cosmo = host.runner_cosmology(cat.cosmology, True)
z_t = np.linspace(0, 30, 1000)
print("D_A table + spline %.2f ms" % best(lambda: interpolate.CubicSpline(z_t, ccl.angular_diameter_distance(cosmo, 1 / (1 + z_t)))))
for nt in (4, 8, 16, 32):
    os.environ["BF_HOST_THREADS"] = str(nt)
    host._pool = None
    print("shell_scalars threads %2d: %.2f ms" % (nt, best(lambda: host.shell_scalars(cat, cat.cosmology, md, 20, model, True))))
os.environ["BF_HOST_THREADS"] = "1"
c = cat.cat
M, z = c["M"], c["z"]
a = 1 / (1 + z)
print("serial pieces (1e6 halos): radius_of %.1f  log_native %.1f  ang2vec %.1f  model_cut %.1f ms" % (
    best(lambda: host.radius_of(md, cosmo, M, a), 2), best(lambda: host.log_native(M), 2),
    best(lambda: host.ang2vec_lonlat(np.asarray(c["ra"]), np.asarray(c["dec"])), 2), best(lambda: host.model_cut(model, M, a), 2)))
