"""Table holders: the duck-type the Runners read from a *model*.

The reference's models expose their tabulated data as plain attributes once
``setup_interpolator`` has run on the host (CCL):

* displacement models (``Profiles/BaryonCorrection.py:306-316``):
  ``raw_input_d`` [Nz,NM,NR], ``raw_input_z_range`` = ln(1+z),
  ``raw_input_M_range`` = ln M, ``raw_input_r_range`` = ln r (or ln r/R_Delta),
  ``Rdelta_sampling``, ``epsilon_max``, ``mass_def``, ``cosmo``, ``p_keys``;
* profile models (``utils/Tabulate.py:229-238``): ``raw_input_2D`` / ``raw_input_3D``
  and the same three axes.

The B200 Runners read exactly those attributes -- from a reference
``Baryonification2D/3D`` / ``TabulatedProfile`` object or from one of the two
holders below (which is what tests and benchmarks use where pyccl is absent).
The holders deliberately have **no** CPU ``displacement()`` / ``projected()``
evaluation: there is no CPU fallback in this package.
"""
from __future__ import annotations

import numpy as np

from ._ccl import default_mass_def


class _NoCpuPath:
    def _nope(self, *a, **k):
        raise RuntimeError("baryonification_b200 has no CPU evaluation path; tables are read by the CUDA kernels")


class DisplacementTable(_NoCpuPath):
    """Holder for a tabulated displacement d(ln(1+z), ln M, ln r) in comoving Mpc."""

    def __init__(self, z, M, r, d, cosmo, epsilon_max, mass_def=None, Rdelta_sampling=False):
        self.raw_input_z_range = np.log(1.0 + np.asarray(z, dtype=np.float64))
        self.raw_input_M_range = np.log(np.asarray(M, dtype=np.float64))
        self.raw_input_r_range = np.log(np.asarray(r, dtype=np.float64))
        self.raw_input_d = np.ascontiguousarray(d, dtype=np.float64)
        assert self.raw_input_d.shape == (self.raw_input_z_range.size, self.raw_input_M_range.size,
                                          self.raw_input_r_range.size)
        self.cosmo = cosmo  # a ccl.Cosmology-like object (used for the model-side R_Delta cut)
        self.epsilon_max = epsilon_max
        self.mass_def = mass_def if mass_def is not None else default_mass_def()
        self.Rdelta_sampling = bool(Rdelta_sampling)
        self.p_keys = []

    displacement = _NoCpuPath._nope


class ProfileTable(_NoCpuPath):
    """Holder for tabulated real-space (3D) and projected (2D) profiles."""

    def __init__(self, z, M, r, prof3D, prof2D, cosmo=None, mass_def=None):
        self.raw_input_z_range = np.log(1.0 + np.asarray(z, dtype=np.float64))
        self.raw_input_M_range = np.log(np.asarray(M, dtype=np.float64))
        self.raw_input_r_range = np.log(np.asarray(r, dtype=np.float64))
        self.raw_input_3D = np.ascontiguousarray(prof3D, dtype=np.float64)
        self.raw_input_2D = np.ascontiguousarray(prof2D, dtype=np.float64)
        self.cosmo = cosmo
        self.mass_def = mass_def if mass_def is not None else default_mass_def()
        self.p_keys = []

    real = projected = _NoCpuPath._nope


class ParamProfileTable(ProfileTable):
    """Holder for profiles tabulated over extra halo properties (the reference's ``ParamTabulatedProfile``,
    utils/Tabulate.py:524-561): ``raw_input_2D/3D`` are [Nz, NM, NR, Np1, ...], ``p_keys`` names the extra axes
    and ``raw_input_<key>_range`` holds each axis; the per-halo values are catalogue columns of the same names."""

    def __init__(self, z, M, r, prof3D, prof2D, other_params, cosmo=None, mass_def=None):
        super().__init__(z, M, r, prof3D, prof2D, cosmo, mass_def)
        self.p_keys = list(other_params.keys())
        for k, v in other_params.items():
            setattr(self, "raw_input_%s_range" % k, np.asarray(v, dtype=np.float64))


class HostTable:
    """Contiguous float64 view of one table, ready for H2D: axes (z, M[, extras...]) + the radial axis.
    ``values`` is stored with r LAST ([Nz, NM, Np1, ..., NR]) so that the per-halo collapse reads rows."""

    __slots__ = ("gz", "gM", "gr", "values", "is_log", "extra", "keys")

    def __init__(self, gz, gM, gr, values, is_log, extra=(), keys=()):
        self.gz = np.ascontiguousarray(gz, dtype=np.float64)
        self.gM = np.ascontiguousarray(gM, dtype=np.float64)
        self.gr = np.ascontiguousarray(gr, dtype=np.float64)
        self.extra = [np.ascontiguousarray(g, dtype=np.float64) for g in extra]
        self.keys = list(keys)
        values = np.asarray(values, dtype=np.float64)
        want = (self.gz.size, self.gM.size, self.gr.size) + tuple(g.size for g in self.extra)
        if values.shape != want:
            raise ValueError(f"table shape {values.shape} does not match its axes {want}")
        if self.extra:  # reference layout [z, M, r, p...] -> [z, M, p..., r]
            values = np.moveaxis(values, 2, -1)
        self.values = np.ascontiguousarray(values)
        self.is_log = bool(is_log)
        for g, nm in [(self.gz, "z"), (self.gM, "M"), (self.gr, "r")] + list(zip(self.extra, self.keys)):
            if g.size < 2 or not np.all(np.diff(g) > 0):
                raise ValueError(f"table axis {nm} must be strictly ascending with >= 2 points")


def _require(model, names):
    missing = [n for n in names if not hasattr(model, n)]
    if missing:
        raise TypeError(
            f"model {type(model).__name__} has no tabulated data ({', '.join(missing)} missing). "
            "Run setup_interpolator() / wrap the profile in TabulatedProfile first: the B200 runners read "
            "tables only and have no CPU fallback for analytic profiles.")


def _no_pkeys(model):
    """The reference's Baryonify runners assert that a model with property keys is a ParamTabulatedProfile
    (HealpixRunner.py:282-287, Map2DRunner.py:469-474) -- which a displacement model never is."""
    keys = list(vars(model).get("p_keys", []) or [])
    assert not keys, (f"You asked to use {keys} properties in Baryonification. You must pass a ParamTabulatedProfile"
                      f"as the model. You have passed {type(model)} instead")


def displacement_host_table(model) -> HostTable:
    """Export ``interp_d``'s grid and values (BaryonCorrection.py:306-315): linear in d."""
    if not hasattr(model, "raw_input_d"):
        if hasattr(model, "displacement") and not hasattr(model, "interp_d"):
            raise NameError("No Table created. Run setup_interpolator() method first")  # BaryonCorrection.py:426
    _require(model, ["raw_input_d", "raw_input_z_range", "raw_input_M_range", "raw_input_r_range"])
    _no_pkeys(model)
    return HostTable(model.raw_input_z_range, model.raw_input_M_range, model.raw_input_r_range,
                     model.raw_input_d, is_log=False)


def profile_host_table(model, which: str) -> HostTable:
    """Export ln(raw_input_2D|3D) exactly as Tabulate.py:237-238 builds it (numpy on host,
    so -inf / NaN entries are bit-identical to the reference's interpolator)."""
    name = "raw_input_2D" if which == "2D" else "raw_input_3D"
    if not hasattr(model, name) and (hasattr(model, "_projected") or hasattr(model, "projected")) \
            and not hasattr(model, "interp2D") and hasattr(model, "setup_interpolator"):
        raise NameError("No Table created. Run setup_interpolator() method first")  # Tabulate.py:322
    keys = list(vars(model).get("p_keys", []) or [])
    _require(model, [name, "raw_input_z_range", "raw_input_M_range", "raw_input_r_range"] +
             ["raw_input_%s_range" % k for k in keys])
    with np.errstate(divide="ignore", invalid="ignore"):
        vals = np.log(np.asarray(getattr(model, name), dtype=np.float64))
    return HostTable(model.raw_input_z_range, model.raw_input_M_range, model.raw_input_r_range, vals, is_log=True,
                     extra=[getattr(model, "raw_input_%s_range" % k) for k in keys], keys=keys)
