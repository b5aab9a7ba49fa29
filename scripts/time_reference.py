#!/usr/bin/env python
"""Build container only (/root/reference present): wall time of the REAL reference Runners (BaryonForge's own Python
loops, executed verbatim through oracle/ref_shim.py: pyccl -> the flat-wCDM stand-in, healpy -> the C restatement) next
to the C/OpenMP oracle port on the SAME inputs, so that the port-vs-reference factor of bench.py's `cpu_baseline`
("kind": "port") is a measured number.  Writes profiles/r02_reference_vs_port.json.

  C1  BaryonifyGrid 2-D 512^2, 2 000 halos, epsilon_max 20 (BASELINE.json configs[0]) -- in full
  C2s BaryonifyShell NSIDE 1024, the first 2 000 halos of the seed-42 shuffle of the 1e5-halo catalogue, epsilon_max 20
"""
import copy
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden")]


def main():
    from oracle import ref_shim
    bfg = ref_shim.import_reference()
    import make_golden as mg
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    ncores = os.cpu_count() or 1
    out = {"host": {"cores": ncores, "container": "build container (no GPU)"}, "shim": ref_shim.describe(), "cases": {}}

    def timeit(fn, reps=1):
        best = None
        for _ in range(reps):
            t0 = time.perf_counter()
            r = fn()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        return best, r

    # ---- C1
    cfg = syn.config_grid()
    with np.errstate(all="ignore"):
        mg.run_reference(bfg, "grid_baryonify", syn.config_grid(npix=64, nhalo=20))  # numba JIT warm-up
        t_ref, ref = timeit(lambda: mg.run_reference(bfg, "grid_baryonify", cfg)["out"])
    r1 = orc.BaryonifyGrid(cfg["catalog"], cfg["field"], 20.0, cfg["model"], nthreads=1)
    t_p1, got = timeit(r1.process, 2)
    rN = orc.BaryonifyGrid(cfg["catalog"], cfg["field"], 20.0, cfg["model"], nthreads=ncores)
    t_pN, _ = timeit(rN.process, 2)
    out["cases"]["C1 BaryonifyGrid 2-D 512^2, 2000 halos"] = {
        "pairs": int(r1.npairs), "reference_python_s": t_ref, "port_1_thread_s": t_p1, f"port_{ncores}_threads_s": t_pN,
        "port_over_reference_1_thread": t_ref / t_p1, f"port_over_reference_{ncores}_threads": t_ref / t_pN,
        "max_rel_diff_port_vs_reference": float(np.max(np.abs(got - ref)) / np.max(np.abs(ref)))}
    # ---- C2 slice
    cfg = syn.config_shell(nside=1024, nhalo=100_000)
    n = len(cfg["catalog"].cat)
    idx = np.sort(np.random.default_rng(42).choice(n, size=n, replace=False)[:2000])
    sub = copy.copy(cfg["catalog"])
    sub.cat = cfg["catalog"].cat[idx]
    cfg["catalog"] = sub
    with np.errstate(all="ignore"):
        mg.run_reference(bfg, "shell_baryonify", syn.config_shell(nside=16, nhalo=20))
        t_ref, ref = timeit(lambda: mg.run_reference(bfg, "shell_baryonify", cfg)["out"])
    empty = copy.copy(sub)
    empty.cat = sub.cat[:0]
    cfg_e = dict(cfg, catalog=empty)
    with np.errstate(all="ignore"):
        t_ref_empty, _ = timeit(lambda: mg.run_reference(bfg, "shell_baryonify", cfg_e)["out"])
    r1 = orc.BaryonifyShell(sub, cfg["field"], 20.0, cfg["model"], nthreads=1)
    t_l1, _ = timeit(r1.offsets)
    t_lN, _ = timeit(orc.BaryonifyShell(sub, cfg["field"], 20.0, cfg["model"], nthreads=ncores).offsets)
    t_p1, got = timeit(r1.process)
    rN = orc.BaryonifyShell(sub, cfg["field"], 20.0, cfg["model"], nthreads=ncores)
    t_pN, _ = timeit(rN.process)
    out["cases"]["C2 slice: BaryonifyShell NSIDE 1024, 2000 of 1e5 halos"] = {
        "pairs": int(r1.npairs), "reference_python_s": t_ref, "port_1_thread_s": t_p1, f"port_{ncores}_threads_s": t_pN,
        "port_over_reference_1_thread": t_ref / t_p1, f"port_over_reference_{ncores}_threads": t_ref / t_pN,
        "max_rel_diff_port_vs_reference": float(np.max(np.abs(got - ref)) / np.max(np.abs(ref))),
        "halo_loop_only": {"reference_python_s": t_ref - t_ref_empty, "reference_pairs_per_s": r1.npairs / (t_ref - t_ref_empty),
                           "reference_halos_per_s": 2000 / (t_ref - t_ref_empty),
                           "port_1_thread_s": t_l1, f"port_{ncores}_threads_s": t_lN,
                           "port_over_reference_1_thread": (t_ref - t_ref_empty) / t_l1,
                           f"port_over_reference_{ncores}_threads": (t_ref - t_ref_empty) / t_lN,
                           "how": "reference: process() minus process() of an empty catalogue (set-up + re-binning); port: "
                                  "offsets() = per-halo scalar loop in Python + the C pair loop"},
        "note": "the whole-process() figures include the O(Npix) re-binning of the 12.6 Mpixel map, serial in both"}
    path = os.path.join(ROOT, "profiles", "r02_reference_vs_port.json")
    json.dump(out, open(path, "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
