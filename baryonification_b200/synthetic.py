"""Seeded synthetic catalogs, fields and tables of the benchmark shapes.

Everything is ``np.random.default_rng(seed)`` (PCG64).  Used by tests, ``bench.py`` and
``__graft_entry__.smoke()``; there is no network, so no real simulation data exists here.
"""
from __future__ import annotations

import numpy as np

from . import _ccl
from .tables import DisplacementTable, ParamProfileTable, ProfileTable
from .utils.io import (GriddedMap, HaloLightConeCatalog, HaloNDCatalog, LightconeShell,
                       ParticleSnapshot)

COSMO = {"Omega_m": 0.3, "Omega_b": 0.045, "h": 0.7, "sigma8": 0.8, "n_s": 0.96, "w0": -1.0}
SEED0 = 20240903


def model_cosmology(cosmo=COSMO):
    ccl = _ccl.get_ccl()
    return ccl.Cosmology(Omega_c=cosmo["Omega_m"] - cosmo["Omega_b"], Omega_b=cosmo["Omega_b"], h=cosmo["h"],
                         sigma8=cosmo["sigma8"], n_s=cosmo["n_s"], w0=cosmo["w0"],
                         matter_power_spectrum="linear")


def halo_masses(rng, n, lo=13.0, hi=15.3, slope=0.9):
    """log10 M by inverse CDF of p ~ 10^(-slope (log10 M - lo)) on [lo, hi]."""
    u = rng.random(n)
    k = slope * np.log(10.0)
    x = -np.log(1.0 - u * (1.0 - np.exp(-k * (hi - lo)))) / k
    return 10.0 ** (lo + x)


def lightcone_halos(n, seed, zmin=0.1, zmax=1.0, cosmo=COSMO):
    rng = np.random.default_rng(seed)
    M = halo_masses(rng, n)
    z = rng.uniform(zmin, zmax, n)
    ra = rng.uniform(0.0, 360.0, n)
    dec = np.degrees(np.arcsin(rng.uniform(-1.0, 1.0, n)))
    return dict(ra=ra, dec=dec, M=M, z=z, cosmo=cosmo)


def box_halos(n, seed, L, dim=3, redshift=0.3, cosmo=COSMO):
    rng = np.random.default_rng(seed)
    M = halo_masses(rng, n)
    xyz = rng.uniform(0.0, L, (n, 3))
    d = dict(x=xyz[:, 0], y=xyz[:, 1], M=M, redshift=redshift, cosmo=cosmo)
    if dim == 3:
        d["z"] = xyz[:, 2]
    return d


def table_axes(zmin, zmax, Mmin=1e13 * 0.9, Mmax=10 ** 15.3 * 1.1, nz=10, nM=10, nR=500, Rmin=1e-3, Rmax=3e2):
    return (np.geomspace(zmin, zmax, nz), np.geomspace(Mmin, Mmax, nM), np.geomspace(Rmin, Rmax, nR))


def _RDelta(M):
    return (M / 1e14) ** (1.0 / 3.0)


def displacement_values(z, M, r):
    """d = 0.05 R x/(1+x^2) tanh(1 - x/3) (1+z): smooth, sign-changing, -> 0."""
    Z, MM, R = np.meshgrid(z, M, r, indexing="ij")
    RD = _RDelta(MM)
    x = R / RD
    return 0.05 * RD * x / (1 + x * x) * np.tanh(1 - x / 3) * (1 + Z)


def profile_values(z, M, r, zero_beyond=None):
    """P = 1e-6 (M/1e14)^(5/3) (1+z)^(8/3) (1+(r/0.2R)^2)^-1.5 exp(-r/4R)."""
    Z, MM, R = np.meshgrid(z, M, r, indexing="ij")
    RD = _RDelta(MM)
    P = 1e-6 * (MM / 1e14) ** (5.0 / 3.0) * (1 + Z) ** (8.0 / 3.0) * (1 + (R / (0.2 * RD)) ** 2) ** -1.5 \
        * np.exp(-R / (4 * RD))
    if zero_beyond is not None:
        P = np.where(R > zero_beyond * RD, 0.0, P)  # exercises the ln 0 = -inf path
    return P


def displacement_model(zmin, zmax, epsilon_max, cosmo=COSMO, nz=10, nM=10, nR=500, cls=DisplacementTable, **kw):
    z, M, r = table_axes(zmin, zmax, nz=nz, nM=nM, nR=nR)
    return cls(z, M, r, displacement_values(z, M, r), model_cosmology(cosmo), epsilon_max, **kw)


def profile_model(zmin, zmax, cosmo=COSMO, nz=10, nM=10, nR=500, zero_beyond=None, cls=ProfileTable):
    z, M, r = table_axes(zmin, zmax, nz=nz, nM=nM, nR=nR)
    P3 = profile_values(z, M, r, zero_beyond)
    P2 = 2.0 * r[None, None, :] * P3 * (1.0 / (1.0 + z))[:, None, None]  # any smooth positive "projection"
    return cls(z, M, r, P3, P2, model_cosmology(cosmo))


CDELTA_GRID = np.linspace(2.0, 12.0, 6)


def param_profile_model(zmin, zmax, cosmo=COSMO, nz=10, nM=10, nR=500, zero_beyond=None):
    """The same profiles tabulated over one extra halo property, ``cdelta`` (p_keys = ['cdelta'], like a
    ParamTabulatedProfile built with other_params = {'cdelta': ...}, utils/Tabulate.py:524-561)."""
    z, M, r = table_axes(zmin, zmax, nz=nz, nM=nM, nR=nR)
    c = CDELTA_GRID
    P3 = profile_values(z, M, r, zero_beyond)[..., None] * ((c / 5.0) ** 1.5 / (1 + (r[:, None] * c / 4.0) ** 0.5))[None, None]
    P2 = 2.0 * r[None, None, :, None] * P3 * (1.0 / (1.0 + z))[:, None, None, None]
    return ParamProfileTable(z, M, r, P3, P2, {"cdelta": c}, model_cosmology(cosmo))


def cdelta_column(n, seed):
    """Per-halo concentrations; a few lie outside the tabulated axis (their readout is NaN -> painted as 0)."""
    rng = np.random.default_rng(seed + 555)
    c = rng.uniform(CDELTA_GRID[0], CDELTA_GRID[-1], n)
    c[::97] = CDELTA_GRID[-1] + 1.0
    c[1::211] = CDELTA_GRID[2]  # exactly on a node
    return c


def lognormal_field(shape, seed):
    rng = np.random.default_rng(seed)
    return rng.lognormal(0.0, 1.0, shape) * 1e10


def grid_bins(npix, L):
    """Pixel-centre coordinates of an ``npix``-per-side mesh over [0, L)."""
    return (np.arange(npix) + 0.5) * (L / npix)


# --------------------------------------------------------------------------------------
# the five BASELINE.json configurations (sizes overridable so tests can shrink them)
# --------------------------------------------------------------------------------------
def config_grid(npix=512, nhalo=2000, dim=2, L=500.0 / 0.7, seed=SEED0 + 1, epsilon_max=20.0, ellipticity=False,
                p_keys=False, rdelta=False):
    halos = box_halos(nhalo, seed, L, dim=dim)
    if p_keys:
        halos["cdelta"] = cdelta_column(nhalo, seed)
    if ellipticity:  # q_ell axis ratios, A_ell major-axis directions (+ c_ell in 3-D), Map2DRunner.py:266-272
        rng = np.random.default_rng(seed + 77)
        halos["q_ell"] = rng.uniform(0.4, 1.0, nhalo)
        halos["q_ell"][:3] = [1.0, 0.99995, 0.5]  # round halo, the small-eta series branch, a flat one
        halos["A_ell"] = rng.normal(size=(nhalo, dim))
        if dim == 3:
            halos["c_ell"] = rng.uniform(0.3, 1.0, nhalo)
    field = lognormal_field((npix,) * dim, seed + 1000)
    cat = HaloNDCatalog(**halos)
    gmap = GriddedMap(map=field, redshift=halos["redshift"], bins=grid_bins(npix, L), cosmo=COSMO)
    z = halos["redshift"]
    model = displacement_model(z * 0.5, z * 2.0, epsilon_max, Rdelta_sampling=rdelta)
    pmodel = param_profile_model(z * 0.5, z * 2.0) if p_keys else profile_model(z * 0.5, z * 2.0)
    return dict(catalog=cat, field=gmap, model=model, paint_model=pmodel, epsilon_max=epsilon_max,
                use_ellipticity=bool(ellipticity))


def config_shell(nside=1024, nhalo=100_000, seed=SEED0 + 2, epsilon_max=20.0, zero_map=False, zero_beyond=None,
                 p_keys=False, rdelta=False):
    halos = lightcone_halos(nhalo, seed)
    if p_keys:
        halos["cdelta"] = cdelta_column(nhalo, seed)
    npix = 12 * nside * nside
    field = np.zeros(npix) if zero_map else lognormal_field(npix, seed + 1000)
    cat = HaloLightConeCatalog(**halos)
    shell = LightconeShell(map=field, cosmo=COSMO)
    model = displacement_model(0.05, 1.2, epsilon_max, Rdelta_sampling=rdelta)
    pmodel = (param_profile_model if p_keys else profile_model)(0.05, 1.2, zero_beyond=zero_beyond)
    return dict(catalog=cat, field=shell, model=model, paint_model=pmodel, epsilon_max=epsilon_max)


def config_snapshot(npart_side=512, nhalo=100_000, dim=3, L=500.0 / 0.7, seed=SEED0 + 4, epsilon_max=5.0, rdelta=False):
    halos = box_halos(nhalo, seed, L, dim=dim)
    rng = np.random.default_rng(seed + 1000)
    n = npart_side ** dim
    xyz = [rng.uniform(0.0, L, n) for _ in range(dim)]
    snap = ParticleSnapshot(x=xyz[0], y=xyz[1], z=xyz[2] if dim == 3 else None, M=1.0, L=L,
                            redshift=halos["redshift"], cosmo=COSMO)
    cat = HaloNDCatalog(**halos)
    z = halos["redshift"]
    model = displacement_model(z * 0.5, z * 2.0, epsilon_max, Rdelta_sampling=rdelta)
    return dict(catalog=cat, field=snap, model=model, epsilon_max=epsilon_max)
