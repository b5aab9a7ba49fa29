"""Shared helpers of the golden-fixture tests: rebuild a case's seeded inputs and run an implementation on it.

`impl` is either the `oracle.runners` module (CPU oracle) or the `baryonification_b200` package (CUDA path);
both expose the reference's five Runner classes with the same constructor.
"""
from __future__ import annotations

import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def meta():
    with open(os.path.join(GOLDEN, "golden.json")) as f:
        return json.load(f)


def case_names(prefix=""):
    return sorted(k for k in meta()["cases"] if k.startswith(prefix))


def build(name):
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(GOLDEN, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    c = meta()["cases"][name]
    cfg = mg.build_case(c["kind"], c["kwargs"])
    assert mg.input_digest(c["kind"], cfg) == c["inputs_sha256"], "seeded inputs drifted from the golden run"
    return c["kind"], cfg


def golden(name):
    with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
        return {k: z[k] for k in z.files}


def run(impl, kind, cfg, **kw):
    """Returns (runner, outputs dict shaped like the fixture)."""
    eps = cfg["epsilon_max"]
    cls = {"shell_baryonify": "BaryonifyShell", "shell_paint": "PaintProfilesShell", "grid_baryonify": "BaryonifyGrid",
           "grid_paint": "PaintProfilesGrid", "snapshot": "BaryonifySnapshot"}[kind]
    model = cfg["paint_model"] if kind.endswith("paint") else cfg["model"]
    if cfg.get("use_ellipticity"):
        kw = dict(kw, use_ellipticity=True)
    r = getattr(impl, cls)(cfg["catalog"], cfg["field"], eps, model, verbose=False, **kw)
    out = r.process()
    if kind == "snapshot":
        dim = 2 if cfg["field"].is2D else 3
        return r, {k: np.asarray(out[k]) for k in ("x", "y", "z")[:dim]}
    return r, {"out": out}


def check(kind, got, want, cfg, rtol=1e-5):
    """north_star tolerances: values within 1e-5 relative (we hold 1e-9 of the map scale), same footprint,
    mass conserved to 1e-10."""
    for k in want:
        g, w = got[k], want[k]
        assert g.shape == w.shape and g.dtype == w.dtype, (k, g.shape, w.shape)
        scale = max(np.max(np.abs(w)), 1e-300)
        assert np.max(np.abs(g - w)) <= 1e-9 * scale, (k, np.max(np.abs(g - w)) / scale)
        if kind.endswith("paint"):
            assert np.array_equal(g != 0, w != 0), "painted footprint differs"
            assert np.allclose(g, w, rtol=rtol, atol=0.0)
    if kind.endswith("baryonify"):
        tot = cfg["field"].map.sum()
        assert abs(got["out"].sum() - tot) / tot < 1e-10
