"""Golden vectors for the Delta P(k) summary, produced by the REFERENCE'S OWN CODE: the code cells of
/root/reference/examples/10_Reproduce_Schneider_deltaPk.ipynb are read from the notebook and executed verbatim
(cell 1: ``numba_histogram3d``; cell 12: the k-bin set-up lines; cell 15: the four fold / histogram / fftn / bincount
lines of the ``for factor`` loop) on small seeded particle sets.  Run in the build container (needs /root/reference):

    python tests/golden/make_golden_pk.py          # writes tests/golden/pk_notebook.npz

``notebook_power()`` is also what tests/test_oracle_pk.py calls live when the reference tree is present.
"""
import gc
import json
import os
import re

import numpy as np

NOTEBOOK = "/root/reference/examples/10_Reproduce_Schneider_deltaPk.ipynb"
HERE = os.path.dirname(os.path.abspath(__file__))

CASES = [  # (name, seed, particles, Ngrd, Nk, L_snap, factor)
    ("ng16_f1", 1, 30_000, 16, 7, 100.0, 1),
    ("ng16_f4", 1, 30_000, 16, 7, 100.0, 4),
    ("ng20_f8", 2, 50_000, 20, 9, 205.0, 8),
    ("ng27_f1", 3, 40_000, 27, 12, 150.0, 1),
]


def notebook_available():
    return os.path.exists(NOTEBOOK)


def _cells():
    nb = json.load(open(NOTEBOOK))
    return ["".join(c["source"]) for c in nb["cells"] if c["cell_type"] == "code"]


def _notebook_sources():
    """(source of numba_histogram3d, the k-bin lines of cell 12, the four P(k) lines of cell 15), verbatim."""
    cells = _cells()
    hist = next(c for c in cells if "def numba_histogram3d" in c)
    setup = next(c for c in cells if re.search(r"^kbins\s*=", c, re.M))
    keep = ("kbins", "klin", "k ", "kinds", "kmsk", "k_c", "k_cen")
    setup_lines = [l for l in setup.splitlines() if re.match(r"^(%s)\s*=" % "|".join(k.strip() for k in keep), l)]
    assert len(setup_lines) == 7, setup_lines
    loop = next(c for c in cells if "MapB = numba_histogram3d(Part_B % Lbox" in c)
    body = [l.strip() for l in loop.splitlines()
            if re.match(r"^\s*(MapB = numba_histogram3d|FFTB = np\.fft\.fftn|FFTB = \(np\.conjugate|PkB  = np\.bincount)", l)]
    assert len(body) == 4, body
    return hist, "\n".join(setup_lines), "\n".join(body)


def particles(seed, n, L):
    """Seeded test particles: uniform + a clustered clump + strays outside [0, L) (the fold wraps them) + exact edges."""
    rng = np.random.default_rng(seed)
    pts = rng.uniform(0.0, L, (n, 3))
    pts[: n // 10] = L * 0.37 + rng.normal(scale=L / 40, size=(n // 10, 3))
    pts[n // 10: n // 10 + 200] += rng.normal(scale=L / 3, size=(200, 3))
    pts[-1] = [0.0, L / 2, L / 4]
    return pts


def notebook_power(points, Ngrd, Nk, L_snap, factor):
    """The notebook's recipe executed from its own source -> dict(Map, Pk, k_c, k_cen)."""
    from numba import njit  # (the notebook's own import)
    hist_src, setup_src, body_src = _notebook_sources()
    ns = {"np": np, "njit": njit, "gc": gc}
    exec(hist_src, ns)
    # cell 12 computes the k bins for Lbox = Snap.L; the factor loop of cell 15 folds into Lbox = Snap.L / factor but
    # keeps the bins and reports k_cen * factor -- the notebook's own convention, reproduced as is
    ns.update(Ngrd=Ngrd, Nk=Nk, Lbox=L_snap)
    with np.errstate(divide="ignore", invalid="ignore"):
        exec(setup_src, ns)
        ns.update(Part_B=np.ascontiguousarray(points), Lbox=L_snap / factor)
        keep_map = body_src.replace("; del MapB; gc.collect()", "")  # (keep the histogram for the fixture)
        exec(keep_map, ns)
    return dict(Map=ns["MapB"], Pk=ns["PkB"], k_c=ns["k_c"], k_cen=ns["k_cen"], kbins=ns["kbins"])


def main():
    out = {}
    for name, seed, n, Ngrd, Nk, L, f in CASES:
        pts = particles(seed, n, L)
        r = notebook_power(pts, Ngrd, Nk, L, f)
        for k, v in r.items():
            out[f"{name}.{k}"] = np.asarray(v)
        out[f"{name}.args"] = np.array([seed, n, Ngrd, Nk, L, f], dtype=np.float64)
    path = os.path.join(HERE, "pk_notebook.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
