"""torchrun --nproc-per-node N scripts/check_sharded.py : sharded runners (NCCL) vs the single-GPU runners."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from oracle import ref_shim

ref_shim.install()
import baryonification_b200 as bf
from baryonification_b200 import sharding as sh, synthetic as syn

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nside = int(sys.argv[1]) if len(sys.argv) > 1 else 512
nh = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
cfg = syn.config_shell(nside=nside, nhalo=nh)
ok = True
for name, cls, model in (("BaryonifyShell", bf.BaryonifyShell, cfg["model"]), ("PaintProfilesShell", bf.PaintProfilesShell, cfg["paint_model"])):
    single = cls(cfg["catalog"], cfg["field"], cfg["epsilon_max"], model, verbose=False, device=local)
    want = single.process()
    sharded = sh.ShardedShell(cls(cfg["catalog"], cfg["field"], cfg["epsilon_max"], model, verbose=False), device=local)
    sharded.process()
    torch.cuda.synchronize(); dist.barrier()
    t0 = time.perf_counter()
    got = sharded.process()
    dt = time.perf_counter() - t0
    same_pairs = sharded.npairs == single.npairs
    err = np.max(np.abs(got - want)) / np.max(np.abs(want))
    exact = np.array_equal(got, want)
    if rank == 0:
        print(f"{name}: world {world} pairs {sharded.npairs} (single {single.npairs}) relerr {err:.2e} bit-identical {exact} "
              f"sharded process() {dt*1e3:.1f} ms", flush=True)
    ok &= same_pairs and err < 1e-12 and (exact or name == "BaryonifyShell")
scfg = syn.config_snapshot(npart_side=128, nhalo=20000, dim=3)
single = bf.BaryonifySnapshot(scfg["catalog"], scfg["field"], scfg["epsilon_max"], scfg["model"], verbose=False, device=local)
want = single.process()
ss = sh.ShardedSnapshot(bf.BaryonifySnapshot(scfg["catalog"], scfg["field"], scfg["epsilon_max"], scfg["model"], verbose=False), device=local)
got = ss.process()
snap_want = want
exact = all(np.array_equal(got[k], want[k]) for k in want.dtype.names)
if rank == 0:
    print(f"BaryonifySnapshot: world {world} pairs {ss.npairs} (single {single.npairs}) bit-identical {exact}", flush=True)
ok &= exact and ss.npairs == single.npairs
for dim, npix, nhalo in ((2, 1000, 5000), (3, 200, 3000)):
    gcfg = syn.config_grid(npix=npix, nhalo=nhalo, dim=dim, epsilon_max=20.0)
    for name, cls, model in (("BaryonifyGrid", bf.BaryonifyGrid, gcfg["model"]), ("PaintProfilesGrid", bf.PaintProfilesGrid, gcfg["paint_model"])):
        single = cls(gcfg["catalog"], gcfg["field"], 20.0, model, verbose=False, device=local)
        want = single.process()
        sg = sh.ShardedGrid(cls(gcfg["catalog"], gcfg["field"], 20.0, model, verbose=False), device=local)
        sg.process()
        torch.cuda.synchronize(); dist.barrier()
        t0 = time.perf_counter()
        got = sg.process()
        dt = time.perf_counter() - t0
        t0 = time.perf_counter(); single.process(); d1 = time.perf_counter() - t0
        exact = np.array_equal(got, want)
        if rank == 0:
            print(f"{name} {dim}-D {npix}: world {world} pairs {sg.npairs} (single {single.npairs}) bit-identical {exact} "
                  f"sharded process() {dt*1e3:.1f} ms (single {d1*1e3:.1f} ms)", flush=True)
        ok &= exact and sg.npairs == single.npairs
# Delta P(k): every rank deposits its slice of the particles, slab-decomposed transform, piecewise k-binning
from baryonification_b200.pk import PowerSpectrumGrid
B = ss.process(gather=False)                                     # this rank's displaced particles
a, b = sh.index_split(len(scfg["field"].cat), world)[rank]
D = scfg["field"].cat[a:b]
for ngrd in (128, 250):
    sp = sh.ShardedPowerSpectrum(ngrd, scfg["field"].L, 60, device=local)
    one = PowerSpectrumGrid(ngrd, scfg["field"].L, 60, device=local)
    res = sp.suppression(B, D, factors=(1, 4))
    ref = one.suppression(snap_want, scfg["field"].cat, factors=(1, 4))
    torch.cuda.synchronize(); dist.barrier()
    t0 = time.perf_counter(); sp.power(B); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    t0 = time.perf_counter(); one.power(one.density(snap_want)); torch.cuda.synchronize(); d1 = time.perf_counter() - t0
    k_s, P_s, n_s = sp.power(B)
    k_1, P_1, n_1 = one.power(one.density(snap_want))
    good = np.array_equal(n_s, n_1)
    m = n_1 > 0
    errP = np.max(np.abs(P_s[m] / P_1[m] - 1))
    errS = max(np.nanmax(np.abs(res[f]["S"] / ref[f]["S"] - 1)) for f in res)
    if rank == 0:
        print(f"DeltaP(k) Ngrd {ngrd}: world {world} mode counts equal {good} P relerr {errP:.2e} S relerr {errS:.2e} "
              f"sharded {dt*1e3:.1f} ms (single {d1*1e3:.1f} ms)", flush=True)
    ok &= good and errP < 1e-10 and errS < 1e-10
dist.barrier()
dist.destroy_process_group()
if rank == 0:
    print("SHARDED CHECK", "OK" if ok else "FAILED", flush=True)
sys.exit(0 if ok else 1)
