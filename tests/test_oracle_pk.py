"""CPU: known answers for the Delta P(k) restatement (oracle/pk.py, the notebook recipe)."""
import numpy as np

from oracle import pk as opk


def test_k_bins_and_mode_counts():
    Ngrd, Nk, L = 16, 7, 100.0
    kinds, kmsk, k_c, k_cen = opk.k_setup(Ngrd, Nk, L)
    kf = 2 * np.pi / L
    assert kinds.size == Ngrd ** 3
    assert k_c[0] >= 6  # the six |k| = k_f modes open the first bin (floor(0) = 0)
    assert k_c.sum() == kmsk.sum() < Ngrd ** 3  # k = 0 and the corners beyond Nyquist are masked
    assert np.all(np.diff(k_cen[k_c > 0]) > 0) and k_cen[0] >= kf


def test_plane_wave_power_lands_in_one_bin():
    Ngrd, Nk, L = 16, 7, 100.0
    g = np.arange(Ngrd)
    # particle counts modulated by cos(2 pi 3 x / L): power only at |k| = 3 k_f
    counts = (10 + 5 * np.cos(2 * np.pi * 3 * g / Ngrd)).round().astype(int)
    pts = []
    for i, c in enumerate(counts):
        for j in range(Ngrd):
            for l in range(Ngrd):
                pts += [[(i + 0.5) * L / Ngrd, (j + 0.5) * L / Ngrd, (l + 0.5) * L / Ngrd]] * int(c)
    pts = np.array(pts)
    k, P, kc, Map = opk.power(pts, Ngrd, Nk, L)
    assert Map.sum() == len(pts) and np.array_equal(Map[:, 0, 0], counts)
    kf = 2 * np.pi / L
    hot = int(np.floor((3 * kf - kf) / ((kf * Ngrd / 2 - kf) / Nk)))
    assert np.argmax(P) == hot and P[hot] > 50 * np.max(np.delete(P, hot))  # (rounding the counts to integers leaks a little power)


def test_folding_wraps_positions():
    rng = np.random.default_rng(2)
    pts = rng.uniform(-20, 130, (5000, 3))
    _, _, _, Map = opk.power(pts, 8, 3, 100.0, factor=4)
    assert Map.sum() == 5000 and Map.min() >= 0


# ------------------------------------------------------------------ pinned against the notebook's own code
def _golden_cases():
    import importlib.util
    import os
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    spec = importlib.util.spec_from_file_location("make_golden_pk", os.path.join(here, "make_golden_pk.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    return mg, np.load(os.path.join(here, "pk_notebook.npz"))


def test_restatement_equals_the_notebooks_own_code_on_the_golden_vectors():
    """tests/golden/pk_notebook.npz was produced by executing the code cells of
    examples/10_Reproduce_Schneider_deltaPk.ipynb verbatim (make_golden_pk.py): histogram, mode counts, P(k) and the
    reported k centres (k_cen * factor, bins of the whole box) of the restatement are the same bits."""
    mg, g = _golden_cases()
    for name, seed, n, Ngrd, Nk, L, f in mg.CASES:
        pts = mg.particles(seed, n, L)
        k, P, kc, Map = opk.power(pts, Ngrd, Nk, L, f)
        assert np.array_equal(Map, g[name + ".Map"]), name
        assert np.array_equal(kc, g[name + ".k_c"]), name
        assert np.array_equal(P, g[name + ".Pk"], equal_nan=True), name
        assert np.array_equal(k, g[name + ".k_cen"] * f, equal_nan=True), name
        assert kc.min() >= 0 and (kc > 0).sum() >= Nk - 1


def test_notebook_code_live_equals_the_golden_vectors():
    """Build container only: the notebook's cells executed now give the committed fixture (the fixture is not stale)."""
    import pytest
    mg, g = _golden_cases()
    if not mg.notebook_available():
        pytest.skip("reference tree not present (GPU box)")
    for name, seed, n, Ngrd, Nk, L, f in mg.CASES[:2]:
        r = mg.notebook_power(mg.particles(seed, n, L), Ngrd, Nk, L, f)
        for key in ("Map", "Pk", "k_c", "k_cen"):
            assert np.array_equal(np.asarray(r[key]), g[f"{name}.{key}"], equal_nan=True), (name, key)
