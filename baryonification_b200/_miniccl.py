"""Minimal stand-in for the handful of pyccl calls the Runner hot path makes.

The reference builds a ``ccl.Cosmology`` in every ``process()`` and uses it for
exactly two per-halo scalars (``Runners/HealpixRunner.py:268-280,296``;
``Runners/Map2DRunner.py:456-460,485``; ``Runners/SnapshotRunner.py:195-199,213``;
``Profiles/BaryonCorrection.py:370``):

* ``mass_def.get_radius(cosmo, M, a)``  -> R_Delta in physical Mpc
* ``ccl.angular_diameter_distance(cosmo, a)`` -> D_A in physical Mpc

When the real ``pyccl`` imports it is used unchanged (see ``_ccl.py``).  This
module exists because pyccl is not installable on the build / GPU boxes (no
network): it implements the same call signatures for a flat wCDM background
without radiation, so that the GPU path, the CPU oracle and the benchmark all
consume *identical* per-halo scalars.  It is host-side input preparation, not
part of the accelerated path, and it is NOT claimed to match CCL to its full
precision (CCL adds photons + neutrinos to E(a)); see DESIGN.md.
"""
from __future__ import annotations

import types

import numpy as np

# CCL's physical constants (pyccl/physical_constants): rho_crit in Msun/h / (Mpc/h)^3
RHO_CRITICAL = 2.7744948e11
CLIGHT_HMPC = 2997.92458  # c / (100 km/s/Mpc) in Mpc/h
_FOUR_THIRDS_PI = 4.18879020479  # literal used by pyccl.halos.massdef.MassDef.get_radius


class Cosmology:
    """Flat wCDM background holder with the ``ccl.Cosmology`` keyword surface."""

    def __init__(self, Omega_c=None, Omega_b=None, h=None, sigma8=None, n_s=None,
                 w0=-1.0, wa=0.0, matter_power_spectrum="linear", **_ignored):
        self._params = dict(Omega_c=float(Omega_c), Omega_b=float(Omega_b), h=float(h),
                            sigma8=sigma8, n_s=n_s, w0=float(w0), wa=float(wa))
        self._chi_tab = None

    def __getitem__(self, key):
        if key == "Omega_m":
            return self._params["Omega_c"] + self._params["Omega_b"]
        return self._params[key]

    # no-ops kept so reference code calling them keeps working
    def compute_sigma(self):
        return None

    def compute_distances(self):
        return None

    def h_over_h0(self, a):
        a = np.asarray(a, dtype=np.float64)
        om = self["Omega_m"]
        w0 = self._params["w0"]
        return np.sqrt(om * a ** -3 + (1.0 - om) * a ** (-3.0 * (1.0 + w0)))

    def rho_x(self, a, species="critical", is_comoving=False):
        """Physical density of ``species`` in Msun / Mpc^3 (CCL convention)."""
        a = np.asarray(a, dtype=np.float64)
        h = self._params["h"]
        rho_c0 = RHO_CRITICAL * h * h
        if species == "critical":
            rho = rho_c0 * self.h_over_h0(a) ** 2
        elif species == "matter":
            rho = rho_c0 * self["Omega_m"] * a ** -3
        else:
            raise ValueError(f"species {species!r} not supported by the pyccl stand-in")
        return rho * a ** 3 if is_comoving else rho

    def comoving_radial_distance(self, a):
        """chi(a) in Mpc by composite Simpson quadrature on a fixed ln(a) grid."""
        if self._chi_tab is None:
            # integrate d chi = c/H0 * da / (a^2 E(a)) from a to 1 on a dense grid
            lna = np.linspace(np.log(1.0 / 31.5), 0.0, 8193)
            a_g = np.exp(lna)
            f = 1.0 / (a_g * self.h_over_h0(a_g))  # integrand in d ln a
            dl = lna[1] - lna[0]
            # cumulative Simpson from the right end (a = 1)
            seg = (f[:-2:2] + 4.0 * f[1:-1:2] + f[2::2]) * dl / 3.0
            cum = np.concatenate([[0.0], np.cumsum(seg[::-1])])[::-1]
            self._chi_tab = (lna[::2], cum * CLIGHT_HMPC / self._params["h"])
        lna_t, chi_t = self._chi_tab
        from scipy.interpolate import CubicSpline
        return CubicSpline(lna_t, chi_t)(np.log(np.asarray(a, dtype=np.float64)))

    def angular_diameter_distance(self, a):
        a = np.asarray(a, dtype=np.float64)
        return a * self.comoving_radial_distance(a)


def angular_diameter_distance(cosmo, a):
    return cosmo.angular_diameter_distance(a)


def comoving_radial_distance(cosmo, a):
    return cosmo.comoving_radial_distance(a)


def h_over_h0(cosmo, a):
    return cosmo.h_over_h0(a)


def rho_x(cosmo, a, species, is_comoving=False):
    return cosmo.rho_x(a, species, is_comoving)


class MassDef:
    """``pyccl.halos.massdef.MassDef`` for fixed-overdensity definitions."""

    def __init__(self, Delta, rho_type, **_ignored):  # c_m_relation etc. are table-construction only
        if isinstance(Delta, str):
            raise ValueError("the pyccl stand-in only supports numeric Delta")
        if rho_type not in ("critical", "matter"):
            raise ValueError("rho_type must be 'critical' or 'matter'")
        self.Delta = Delta
        self.rho_type = rho_type

    def get_Delta(self, cosmo, a):
        return self.Delta

    def get_radius(self, cosmo, M, a):
        M_use = np.atleast_1d(M)
        Delta = self.get_Delta(cosmo, a)
        R = (M_use / (_FOUR_THIRDS_PI * Delta * cosmo.rho_x(a, self.rho_type))) ** (1.0 / 3.0)
        if np.ndim(M) == 0:
            return R[0]
        return R

    def get_mass(self, cosmo, R, a):
        R_use = np.atleast_1d(R)
        Delta = self.get_Delta(cosmo, a)
        M = 4.18879020479 * cosmo.rho_x(a, self.rho_type) * Delta * R_use ** 3
        if np.ndim(R) == 0:
            return M[0]
        return M

    def __repr__(self):
        return f"MassDef(Delta={self.Delta}, rho_type={self.rho_type})"

    def __eq__(self, other):
        return (isinstance(other, MassDef) and self.Delta == other.Delta
                and self.rho_type == other.rho_type)

    def __hash__(self):
        return hash((self.Delta, self.rho_type))


class HaloProfile:
    """Base class with the ``real`` / ``projected`` -> ``_real`` / ``_projected`` dispatch."""

    def __init__(self, *args, **kwargs):
        pass

    def real(self, cosmo, r, M, a, *args, **kwargs):
        return self._real(cosmo, r, M, a, *args, **kwargs)

    def projected(self, cosmo, r, M, a, *args, **kwargs):
        return self._projected(cosmo, r, M, a, *args, **kwargs)


def as_module(name="pyccl"):
    """Package the stand-in as a module tree shaped like pyccl (for sys.modules)."""
    top = types.ModuleType(name)
    top.__dict__.update(Cosmology=Cosmology, angular_diameter_distance=angular_diameter_distance,
                        comoving_radial_distance=comoving_radial_distance, h_over_h0=h_over_h0,
                        rho_x=rho_x, __version__="0+miniccl", __miniccl__=True)
    halos = types.ModuleType(name + ".halos")
    massdef = types.ModuleType(name + ".halos.massdef")
    massdef.MassDef = MassDef
    profiles = types.ModuleType(name + ".halos.profiles")
    profiles.HaloProfile = HaloProfile
    halos.massdef, halos.profiles, halos.MassDef, halos.HaloProfile = massdef, profiles, MassDef, HaloProfile
    pc = types.ModuleType(name + ".physical_constants")
    pc.__dict__.update(RHO_CRITICAL=RHO_CRITICAL, CLIGHT_HMPC=CLIGHT_HMPC, CLIGHT=299792458.0,
                       GNEWT=6.67408e-11, SOLAR_MASS=1.988409870698051e+30, MPC_TO_METER=3.085677581491367e+22,
                       PC_TO_METER=3.085677581491367e+16, KBOLTZ=1.38064852e-23, STBOLTZ=5.670367e-08,
                       HPLANCK=6.62607004e-34, EV_IN_J=1.6021766208e-19, T_CMB=2.725)
    conc = types.ModuleType(name + ".halos.concentration")

    class _Concentration:  # placeholder: concentrations are only used in (host) table construction
        def __init__(self, *a, **k):
            pass

    def _conc_getattr(attr):
        if attr.startswith("Concentration"):
            return _Concentration
        raise AttributeError(attr)
    conc.__getattr__ = _conc_getattr
    halos.concentration = conc
    pyutils = types.ModuleType(name + ".pyutils")
    pyutils._fftlog_transform = None
    top.halos, top.physical_constants, top.pyutils = halos, pc, pyutils
    top.__path__ = []  # mark as package
    mods = {name: top, name + ".halos": halos, name + ".halos.massdef": massdef,
            name + ".halos.profiles": profiles, name + ".halos.concentration": conc, name + ".physical_constants": pc,
            name + ".pyutils": pyutils}
    return mods
