#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference (/root/reference/BaryonForge) here.

Build-container only: the reference tree does not exist on the GPU box, so its outputs travel as these small
fixtures.  The reference's two missing third-party dependencies are supplied by oracle/ref_shim.py (pyccl ->
baryonification_b200._miniccl closed forms; healpy -> oracle/healpix.py, the C restatement of the RING routines).
Every per-halo loop, the scipy RegularGridInterpolator readouts, the numba re-binning kernels and the KD-tree ball
queries below are the reference's own code, executed verbatim.

    python tests/golden/make_golden.py          # rewrites tests/golden/*.npz

Inputs are NOT stored: each case is rebuilt from baryonification_b200.synthetic with the recorded keyword
arguments (seeded PCG64), so a fixture holds only the reference's outputs (+ a checksum of the inputs).
"""
from __future__ import annotations

import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.dirname(HERE)):
    if p not in sys.path:
        sys.path.insert(0, p)

# (name, kind, builder kwargs, runner kwargs)
CASES = [
    ("shell_baryonify_ns16", "shell_baryonify", dict(nside=16, nhalo=150, epsilon_max=20.0), {}),
    ("shell_baryonify_ns32_eps4", "shell_baryonify", dict(nside=32, nhalo=300, epsilon_max=4.0), {}),
    ("shell_baryonify_tiny_ns8", "shell_baryonify", dict(nside=8, nhalo=200, epsilon_max=1.0), {}),
    ("shell_baryonify_ns64", "shell_baryonify", dict(nside=64, nhalo=400, epsilon_max=20.0), {}),
    ("shell_paint_ns64", "shell_paint", dict(nside=64, nhalo=400, epsilon_max=20.0, zero_map=True), {}),
    ("shell_paint_ns128_zeros", "shell_paint", dict(nside=128, nhalo=300, epsilon_max=30.0, zero_map=True,
                                                    zero_beyond=50.0), {}),
    ("grid2d_baryonify_64", "grid_baryonify", dict(npix=64, nhalo=120, dim=2, epsilon_max=20.0), {}),
    ("grid3d_baryonify_24", "grid_baryonify", dict(npix=24, nhalo=40, dim=3, epsilon_max=8.0, L=100.0), {}),
    ("grid2d_paint_64", "grid_paint", dict(npix=64, nhalo=120, dim=2, epsilon_max=20.0), {}),
    ("grid3d_paint_24", "grid_paint", dict(npix=24, nhalo=40, dim=3, epsilon_max=8.0, L=100.0), {}),
    ("grid2d_baryonify_ell_64", "grid_baryonify", dict(npix=64, nhalo=120, dim=2, epsilon_max=20.0, ellipticity=True), {}),
    ("grid2d_paint_ell_64", "grid_paint", dict(npix=64, nhalo=120, dim=2, epsilon_max=20.0, ellipticity=True), {}),
    ("shell_paint_pkeys_ns32", "shell_paint", dict(nside=32, nhalo=300, epsilon_max=20.0, zero_map=True, p_keys=True), {}),
    ("grid2d_paint_pkeys_64", "grid_paint", dict(npix=64, nhalo=120, dim=2, epsilon_max=20.0, p_keys=True), {}),
    ("shell_baryonify_rdelta_ns32", "shell_baryonify", dict(nside=32, nhalo=300, epsilon_max=20.0, rdelta=True), {}),
    ("grid3d_baryonify_rdelta_24", "grid_baryonify", dict(npix=24, nhalo=40, dim=3, epsilon_max=8.0, L=100.0, rdelta=True), {}),
    ("snapshot3d_rdelta_16", "snapshot", dict(npart_side=16, nhalo=80, dim=3, epsilon_max=5.0, L=100.0, rdelta=True), {}),
    ("snapshot3d_16", "snapshot", dict(npart_side=16, nhalo=80, dim=3, epsilon_max=5.0, L=100.0), {}),
    ("snapshot2d_64", "snapshot", dict(npart_side=64, nhalo=60, dim=2, epsilon_max=5.0, L=300.0), {}),
]


def build_case(kind, kw):
    from baryonification_b200 import synthetic as syn
    if kind.startswith("shell"):
        return syn.config_shell(**kw)
    if kind.startswith("grid"):
        return syn.config_grid(**kw)
    return syn.config_snapshot(**kw)


def input_digest(kind, cfg):
    h = hashlib.sha256()
    cat = cfg["catalog"].cat
    for k in cat.dtype.names:
        h.update(np.ascontiguousarray(cat[k]).tobytes())
    f = cfg["field"]
    if hasattr(f, "map"):
        h.update(np.ascontiguousarray(f.map).tobytes())
    else:
        for k in f.cat.dtype.names:
            h.update(np.ascontiguousarray(f.cat[k]).tobytes())
    m = cfg["paint_model"] if kind.endswith("paint") else cfg["model"]
    for k in ("raw_input_d", "raw_input_2D", "raw_input_3D"):
        if hasattr(m, k):
            h.update(np.ascontiguousarray(getattr(m, k)).tobytes())
    return h.hexdigest()


def run_reference(bfg, kind, cfg):
    """Reference Runner objects fed with the reference's own container / model classes."""
    import refmodels as rm
    eps = cfg["epsilon_max"]
    if kind == "shell_baryonify":
        r = bfg.BaryonifyShell(rm.ref_lightcone_catalog(bfg, cfg["catalog"]), rm.ref_shell(bfg, cfg["field"]), eps,
                               rm.ref_displacement_model(bfg, cfg["model"]), verbose=False)
        return {"out": r.process()}
    if kind == "shell_paint":
        r = bfg.PaintProfilesShell(rm.ref_lightcone_catalog(bfg, cfg["catalog"]), rm.ref_shell(bfg, cfg["field"]), eps,
                                   rm.ref_profile_model(bfg, cfg["paint_model"]), verbose=False)
        return {"out": r.process()}
    dim = 2 if getattr(cfg["field"], "is2D", False) else 3
    if kind == "grid_baryonify":
        r = bfg.BaryonifyGrid(rm.ref_nd_catalog(bfg, cfg["catalog"], dim), rm.ref_grid(bfg, cfg["field"]), eps,
                              rm.ref_displacement_model(bfg, cfg["model"]), verbose=False,
                              use_ellipticity=cfg.get("use_ellipticity", False))
        return {"out": r.process()}
    if kind == "grid_paint":
        r = bfg.PaintProfilesGrid(rm.ref_nd_catalog(bfg, cfg["catalog"], dim), rm.ref_grid(bfg, cfg["field"]), eps,
                                  rm.ref_profile_model(bfg, cfg["paint_model"]), verbose=False,
                                  use_ellipticity=cfg.get("use_ellipticity", False))
        return {"out": r.process()}
    if kind == "snapshot":
        r = bfg.BaryonifySnapshot(rm.ref_nd_catalog(bfg, cfg["catalog"], dim), rm.ref_snapshot(bfg, cfg["field"]), eps,
                                  rm.ref_displacement_model(bfg, cfg["model"]), verbose=False)
        new = r.process()
        return {k: np.asarray(new[k]) for k in ("x", "y", "z")[:dim]}
    raise ValueError(kind)


def readout_vectors(bfg):
    """model.displacement / model.projected of the reference (scipy RGI) at probe radii, incl. out-of-table."""
    import refmodels as rm
    from baryonification_b200 import synthetic as syn
    dm = syn.displacement_model(0.05, 1.2, 20.0)
    pm = syn.profile_model(0.05, 1.2, zero_beyond=50.0)
    rd, rp = rm.ref_displacement_model(bfg, dm), rm.ref_profile_model(bfg, pm)
    r = np.concatenate([np.geomspace(5e-4, 5e2, 400), np.exp(dm.raw_input_r_range[[0, 1, -2, -1]])])
    Ms = np.array([3e13, 1e15, 2e12, 1e14, float(np.exp(dm.raw_input_M_range[3]))])
    zs = np.array([0.3, 0.9, 0.3, 1.5, 0.2])
    d, p2, p3 = [], [], []
    with np.errstate(all="ignore"):
        for M, z in zip(Ms, zs):
            a = 1 / (1 + z)
            d.append(rd.displacement(r, M, a))
            p2.append(rp.projected(None, r, M, a))
            p3.append(rp.real(None, r, M, a))
    return dict(r=r, M=Ms, z=zs, displacement=np.array(d), projected=np.array(p2), real=np.array(p3))


def main():
    from oracle import ref_shim
    shim = ref_shim.install()
    bfg = ref_shim.import_reference()
    meta = {"shim": shim, "numpy": np.__version__, "cases": {}}
    for name, kind, kw, _ in CASES:
        cfg = build_case(kind, kw)
        with np.errstate(all="ignore"):
            out = run_reference(bfg, kind, cfg)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        meta["cases"][name] = {"kind": kind, "kwargs": kw, "inputs_sha256": input_digest(kind, cfg)}
        print(name, {k: v.shape for k, v in out.items()})
    np.savez_compressed(os.path.join(HERE, "readout.npz"), **readout_vectors(bfg))
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
