"""oracle/pk.py -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the Delta P(k) recipe of the reference's notebook
examples/10_Reproduce_Schneider_deltaPk.ipynb (cell 1: numba_histogram3d; cell 12: k bins, computed once for the whole
box; cells 15-21: folding, fftn, bincount, k_cen * factor), statement by statement in numpy.  The notebook is the only
place the reference has this code.  PINNED: tests/golden/make_golden_pk.py executes the notebook's own cells verbatim
on seeded particle sets (tests/golden/pk_notebook.npz); tests/test_oracle_pk.py holds this restatement to those outputs
bit for bit (histograms, mode counts, P(k), k centres) and re-runs the notebook code live when /root/reference is there.
"""
import numpy as np


def histogram3d(points, bins, min_vals, max_vals):  # cell 1 (numba_histogram3d), vectorised
    bin_width = (max_vals - min_vals) / bins
    idx = ((points - min_vals) / bin_width).astype(np.int64)
    idx = np.clip(idx, 0, bins - 1)
    counts = np.zeros((bins, bins, bins), dtype=np.int64)
    np.add.at(counts, (idx[:, 0], idx[:, 1], idx[:, 2]), 1)
    return counts


def k_setup(Ngrd, Nk, Lbox):  # cell 12
    kbins = np.linspace(2 * np.pi / Lbox, 2 * np.pi / Lbox * Ngrd / 2, Nk + 1)
    klin = np.fft.fftfreq(Ngrd, 1 / (2 * np.pi / (Lbox)) / Ngrd)
    k = np.sqrt(klin[:, None, None] ** 2 + klin[None, None, :] ** 2 + klin[None, :, None] ** 2).flatten()
    kinds = np.floor((k - kbins[0]) / (kbins[1] - kbins[0])).astype(int)
    kmsk = (kinds >= 0) & (kinds < Nk)
    k_c = np.bincount(kinds[kmsk], minlength=Nk)
    with np.errstate(divide="ignore", invalid="ignore"):
        k_cen = np.bincount(kinds[kmsk], minlength=Nk, weights=k[kmsk]) / k_c
    return kinds, kmsk, k_c, k_cen


def power(points, Ngrd, Nk, Lsnap, factor=1):  # cells 15-21, one particle set
    kinds, kmsk, k_c, k_cen = k_setup(Ngrd, Nk, Lsnap)  # cell 12: the bins of the WHOLE box, whatever the folding factor
    Lbox = Lsnap / factor
    Map = histogram3d(points % Lbox, bins=Ngrd, min_vals=0.0, max_vals=Lbox)
    FFT = np.fft.fftn(Map)
    FFT = (np.conjugate(FFT) * FFT).real.flatten()
    with np.errstate(divide="ignore", invalid="ignore"):
        Pk = np.bincount(kinds[kmsk], minlength=Nk, weights=FFT[kmsk]) / k_c
    return k_cen * factor, Pk, k_c, Map
