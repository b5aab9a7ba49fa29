"""Stage timings of the shell engines (device-resident inputs).  Usage: perf_shell.py NSIDE NHALO [mode] [variants]"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from baryonification_b200 import synthetic as syn, engine, host, _device as dv
from baryonification_b200.tables import displacement_host_table, profile_host_table
from baryonification_b200._ccl import default_mass_def

nside, nh = int(sys.argv[1]), int(sys.argv[2])
mode = sys.argv[3] if len(sys.argv) > 3 else "baryonify"
variants = sys.argv[4].split(",") if len(sys.argv) > 4 else ["gather", "atomic"]
cfg = syn.config_shell(nside=nside, nhalo=nh)
m = engine.BARYONIFY if mode == "baryonify" else engine.PAINT
t0 = time.time()
model = cfg["model"] if m else cfg["paint_model"]
s = host.shell_scalars(cfg["catalog"], cfg["catalog"].cosmo, default_mass_def(), cfg["epsilon_max"], model, bool(m))
htab = displacement_host_table(model) if m else profile_host_table(model, "2D")
print(f"host setup {time.time()-t0:.3f}s")
dev = dv.device(0)
orig = dv.to_dev(cfg["field"].map)
def ev(fn):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record(); r = fn(); b.record(); torch.cuda.synchronize()
    return r, a.elapsed_time(b)
res = {}
for var in variants:
    eng = engine.ShellEngine(nside, m, htab, s, dev, variant=var)
    _, t_up = ev(lambda: eng.upload())
    _, t_bin = ev(lambda: eng.bin())
    out, t_acc0 = ev(lambda: eng.accumulate())
    out, t_acc = ev(lambda: eng.accumulate(out))
    out, t_acc2 = ev(lambda: eng.accumulate(out))
    npairs = eng.pair_count()
    line = f"{var:7s} upload {t_up:8.2f} ms  bin {t_bin:8.2f} ms (entries {eng.nentries})  accumulate {t_acc0:9.2f}/{t_acc:9.2f}/{t_acc2:9.2f} ms  pairs {npairs:.4e}  -> {npairs/t_acc2*1e-6:9.2f} Gpairs/s"
    if m:
        new, t_rg = ev(lambda: eng.regrid(orig, out))
        for meth in ("gather", "sorted", "atomic"):
            new, t_rg = ev(lambda: eng.regrid(orig, out, new, method=meth))
            line += f"  regrid[{eng.regrid_used}] {t_rg:8.2f} ms"
        line += f" window {getattr(eng, 'regrid_window', None)} mass err {abs(new.sum().item()/orig.sum().item()-1):.2e}"
    print(line, flush=True)
    res[var] = out.clone()
if len(res) == 2:
    a, b = res[variants[0]], res[variants[1]]
    print("variants max abs diff", (a - b).abs().max().item(), "max abs", b.abs().max().item())
