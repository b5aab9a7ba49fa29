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
