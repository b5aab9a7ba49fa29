"""Build *reference* (BaryonForge) model / container objects from the same synthetic arrays the
B200 package uses.  Build-container only (needs /root/reference); used by the golden-vector
generator and by tests/test_oracle_vs_reference.py.
"""
from __future__ import annotations

import numpy as np
from scipy import interpolate


def ref_displacement_model(bfg, table):
    """A reference ``Baryonification2D`` carrying ``table``'s (DisplacementTable) data, assembled the
    way ``setup_interpolator`` leaves it (Profiles/BaryonCorrection.py:306-316) without running CCL."""
    m = object.__new__(bfg.Baryonification2D)
    grid = (table.raw_input_z_range, table.raw_input_M_range, table.raw_input_r_range)
    m.raw_input_d = table.raw_input_d
    m.raw_input_z_range, m.raw_input_M_range, m.raw_input_r_range = grid
    m.interp_d = interpolate.RegularGridInterpolator(grid, table.raw_input_d, bounds_error=False, fill_value=np.nan)
    m.Rdelta_sampling = table.Rdelta_sampling
    m.p_keys = []
    m.cosmo, m.epsilon_max, m.mass_def = table.cosmo, table.epsilon_max, table.mass_def
    return m


def ref_profile_model(bfg, table):
    """A reference ``TabulatedProfile`` as ``setup_interpolator`` leaves it (utils/Tabulate.py:229-238), or a
    ``ParamTabulatedProfile`` (:553-561) when the table has property axes."""
    keys = list(getattr(table, "p_keys", []) or [])
    m = object.__new__(bfg.utils.ParamTabulatedProfile if keys else bfg.TabulatedProfile)
    grid = (table.raw_input_z_range, table.raw_input_M_range, table.raw_input_r_range)
    if keys:
        m.p_keys = keys
        for k in keys:
            setattr(m, "raw_input_%s_range" % k, getattr(table, "raw_input_%s_range" % k))
        grid = grid + tuple(getattr(table, "raw_input_%s_range" % k) for k in keys)
    m.raw_input_3D, m.raw_input_2D = table.raw_input_3D, table.raw_input_2D
    m.raw_input_z_range, m.raw_input_M_range, m.raw_input_r_range = grid[:3]
    with np.errstate(divide="ignore", invalid="ignore"):
        m.interp3D = interpolate.RegularGridInterpolator(grid, np.log(table.raw_input_3D), bounds_error=False)
        m.interp2D = interpolate.RegularGridInterpolator(grid, np.log(table.raw_input_2D), bounds_error=False)
    m.cosmo, m.mass_def, m.model = table.cosmo, table.mass_def, None
    return m


def ref_lightcone_catalog(bfg, cat):
    c = cat.cat
    extra = {k: c[k] for k in c.dtype.names if k not in ("M", "z", "ra", "dec")}  # cdelta, ...
    return bfg.HaloLightConeCatalog(ra=c["ra"], dec=c["dec"], M=c["M"], z=c["z"], cosmo=cat.cosmo, **extra)


def ref_nd_catalog(bfg, cat, dim):
    c = cat.cat
    extra = {k: c[k] for k in c.dtype.names if k not in ("M", "x", "y", "z")}  # q_ell, A_ell, ...
    return bfg.HaloNDCatalog(x=c["x"], y=c["y"], z=c["z"] if dim == 3 else None, M=c["M"],
                             redshift=cat.redshift, cosmo=cat.cosmo, **extra)


def ref_shell(bfg, shell):
    return bfg.LightconeShell(map=shell.map, cosmo=shell.cosmo)


def ref_grid(bfg, g):
    return bfg.GriddedMap(map=g.map, redshift=g.redshift, bins=g.bins, cosmo=g.cosmo)


def ref_snapshot(bfg, s):
    c = s.cat
    return bfg.ParticleSnapshot(x=c["x"], y=c["y"], z=None if s.is2D else c["z"], M=c["M"], L=s.L,
                                redshift=s.redshift, cosmo=s.cosmo)
