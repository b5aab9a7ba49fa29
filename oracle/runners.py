"""oracle/runners.py -- TEST INFRASTRUCTURE ONLY (CPU oracle, Python glue).

Restates the five reference Runner ``process()`` methods on top of the C loops in
``runners_oracle.c``.  The per-halo scalar set-up is done here, halo by halo, the
way the reference does it (file:line cited per function); the hot per-pair body
lives in C so the oracle finishes in seconds at test sizes and can serve as the
multi-threaded CPU baseline in ``bench.py``.

Pinned by: tests/test_oracle_vs_reference.py (reference run verbatim through
oracle/ref_shim.py, build container only) and tests/golden/*.npz (outputs of
that run, committed).  HEALPix geometry: parity unpinned (see healpix_oracle.c).

Nothing under ``baryonification_b200/`` may import this module.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
from scipy import interpolate

from . import healpix as ohp

_i64 = C.c_int64
_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_TAB = [_dp, _i64, _dp, _i64, _dp, _i64, _dp]
_ready = False


def _lib():
    global _ready
    L = ohp.lib()
    if not _ready:
        L.bfo_readout.argtypes = _TAB + [C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, _i64, _dp, _dp]
        L.bfo_paint_shell.restype = _i64
        L.bfo_paint_shell.argtypes = [_i64, _i64, _dp, _dp, _dp, _dp, _dp, _dp] + _TAB + [_dp, C.c_int]
        L.bfo_baryonify_shell_offsets.restype = _i64
        L.bfo_baryonify_shell_offsets.argtypes = ([_i64, _i64] + [_dp] * 10 + _TAB + [_dp, C.c_int])
        L.bfo_regrid_shell.argtypes = [_i64, _dp, _dp, _dp]
        L.bfo_regrid_shell_range.argtypes = [_i64, _i64, _i64, _dp, _dp, _dp]
        L.bfo_grid_halos.restype = _i64
        L.bfo_grid_halos.argtypes = ([C.c_int, C.c_int, _i64, _dp, C.c_double, _i64, _dp, _ip, _dp, _dp, _dp, _dp,
                                      C.c_void_p] + _TAB + [_dp, C.c_int])
        L.bfo_grid_regrid.argtypes = [C.c_int, _i64, _dp, _dp, _dp]
        L.bfo_snapshot_offsets.restype = _i64
        L.bfo_snapshot_offsets.argtypes = ([C.c_int, _i64, _dp, C.c_double, _i64, _dp, _dp, C.c_double,
                                            _dp, _dp, _dp, _dp] + _TAB + [_dp, C.c_int])
        L.bfo_snapshot_apply.argtypes = [_i64, _dp, _dp, C.c_double, _dp]
        L.bfo_snapshot_pair_list.restype = _i64
        L.bfo_snapshot_pair_list.argtypes = [C.c_int, _i64, _dp, C.c_double, _i64, _dp, _dp,
                                             np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS"), _i64, C.c_int]
        L.bfo_set_extra_axes.argtypes = [C.c_int, _dp, _ip, _dp]
        _ready = True
    return L


def _c(a, dt=np.float64):
    return np.ascontiguousarray(a, dtype=dt)


# ----------------------------------------------------------------------------
# table adapters
# ----------------------------------------------------------------------------
class _Table:
    def __init__(self, gz, gM, gr, tab):
        self.gz, self.gM, self.gr, self.tab = _c(gz), _c(gM), _c(gr), _c(tab)
        assert self.tab.shape[:3] == (self.gz.size, self.gM.size, self.gr.size)  # (+ p_keys axes after r)

    def args(self):
        return [self.gz, self.gz.size, self.gM, self.gM.size, self.gr, self.gr.size, self.tab]


def displacement_table(model):
    """interp_d grid/values as built at Profiles/BaryonCorrection.py:306-315."""
    if len(getattr(model, "p_keys", [])) > 0:
        raise NotImplementedError("oracle: p_keys tables not restated yet")
    return _Table(model.raw_input_z_range, model.raw_input_M_range, model.raw_input_r_range,
                  model.raw_input_d)


class extra_axes:
    """Context manager: registers the model's extra property axes (``p_keys``: Tabulate.py:524-561 /
    BaryonCorrection.py:205-221) and the per-halo property values ``o_j = {key: cat[key][j]}``
    (HealpixRunner.py:425, Map2DRunner.py:720) with the C oracle for the duration of one runner call."""

    def __init__(self, model, cat):
        keys = list(getattr(model, "p_keys", []) or [])
        self.nx = len(keys)
        if self.nx:
            grids = [_c(getattr(model, "raw_input_%s_range" % k)) for k in keys]
            self.grids = _c(np.concatenate(grids))
            self.sizes = np.array([g.size for g in grids], dtype=np.int64)
            # kwargs[k] * empty (Tabulate.py:607): the catalogue scalar (float32 for ND catalogues) widened to float64
            self.xin = _c(np.stack([np.asarray(cat[k]).astype(np.float64) for k in keys], axis=1))

    def __enter__(self):
        if self.nx:
            _lib().bfo_set_extra_axes(self.nx, self.grids, self.sizes, self.xin)
        return self

    def __exit__(self, *exc):
        if self.nx:
            _lib().bfo_set_extra_axes(0, self.grids, self.sizes, self.xin)
        return False


def profile_table(model, which):
    """log(raw_input_2D|3D) as built at utils/Tabulate.py:229-238 (:553-561 with p_keys axes)."""
    raw = model.raw_input_2D if which == "2D" else model.raw_input_3D
    with np.errstate(divide="ignore", invalid="ignore"):
        return _Table(model.raw_input_z_range, model.raw_input_M_range, model.raw_input_r_range,
                      np.log(raw))


def readout_displacement(model, r, M, a):
    """BaryonificationClass._readout for scalar M, a (BaryonCorrection.py:356-390)."""
    T = displacement_table(model)
    r = _c(np.atleast_1d(r))
    R = model.mass_def.get_radius(model.cosmo, M, a) / a
    shift = float(np.log(R)) if model.Rdelta_sampling else 0.0
    out = np.empty_like(r)
    _lib().bfo_readout(*T.args(), float(np.log(1 / a)), float(np.log(M)), shift,
                       float(model.epsilon_max * R), 0, r.size, r, out)
    return out


def readout_profile(model, r, M, a, which="2D"):
    """TabulatedProfile._readout for scalar M, a (Tabulate.py:272-294)."""
    T = profile_table(model, which)
    r = _c(np.atleast_1d(r))
    out = np.empty_like(r)
    _lib().bfo_readout(*T.args(), float(np.log(1 / a)), float(np.log(M)), 0.0, 0.0, 1, r.size, r, out)
    return out


def _ccl():
    import pyccl as ccl  # real pyccl, or the stand-in installed by oracle.ref_shim
    return ccl


def _log_native(M):
    """``np.log(M_use[i])`` of BaryonCorrection.py:369 / Tabulate.py:283, halo by halo, in the
    catalogue's own dtype (float32 for HaloNDCatalog, io.py:204), then widened to float64."""
    return _c([np.log(M[j]) for j in range(len(M))])


def _model_cut(model, M, a):
    """eps_model * R_model and the Rdelta shift, per halo (BaryonCorrection.py:370-381)."""
    n = len(M)
    cut, shift = np.empty(n), np.zeros(n)
    for j in range(n):
        R = model.mass_def.get_radius(model.cosmo, M[j], a[j]) / a[j]
        cut[j] = model.epsilon_max * R
        if model.Rdelta_sampling:
            shift[j] = np.log(R)
    return cut, shift


# ----------------------------------------------------------------------------
# Shell runners (Runners/HealpixRunner.py)
# ----------------------------------------------------------------------------
class _ShellBase:
    def __init__(self, HaloLightConeCatalog, LightconeShell, epsilon_max, model, use_ellipticity=False,
                 mass_def=None, verbose=True, nthreads=1):
        self.HaloLightConeCatalog, self.LightconeShell = HaloLightConeCatalog, LightconeShell
        self.cosmo = HaloLightConeCatalog.cosmology
        self.model, self.epsilon_max = model, epsilon_max
        self.mass_def = mass_def if mass_def is not None else _ccl().halos.massdef.MassDef(200, "critical")
        self.verbose, self.use_ellipticity, self.nthreads = verbose, use_ellipticity, nthreads
        self.npairs = None
        if use_ellipticity:  # HealpixRunner.py:164-165
            raise NotImplementedError("You have set use_ellipticity = True, but this not yet implemented for HealpixRunner")

    def _halo_scalars(self):
        """HealpixRunner.py:268-305 (and :392-431), halo by halo."""
        ccl = _ccl()
        c = self.cosmo
        cosmo = ccl.Cosmology(Omega_c=c["Omega_m"] - c["Omega_b"], Omega_b=c["Omega_b"], h=c["h"],
                              sigma8=c["sigma8"], n_s=c["n_s"], w0=c["w0"], matter_power_spectrum="linear")
        z_t = np.linspace(0, 30, 1000)
        D_a = interpolate.CubicSpline(z_t, ccl.angular_diameter_distance(cosmo, 1 / (1 + z_t)))
        cat = self.HaloLightConeCatalog.cat
        n = cat.size
        out = {k: np.empty(n) for k in ("M", "a", "R", "D", "radius", "z_in", "M_in")}
        for j in range(n):
            M_j, z_j = cat["M"][j], cat["z"][j]
            a_j = 1 / (1 + z_j)
            R_j = self.mass_def.get_radius(cosmo, M_j, a_j)
            D_j = D_a(z_j)
            out["M"][j], out["a"][j], out["R"][j], out["D"][j] = M_j, a_j, R_j, D_j
            out["radius"][j] = R_j * self.epsilon_max / D_j
            out["z_in"][j] = np.log(1 / a_j)
            out["M_in"][j] = np.log(M_j)
        out["lon"], out["lat"] = _c(cat["ra"]), _c(cat["dec"])
        out["vec"] = _c(ohp.ang2vec(out["lon"], out["lat"], lonlat=True)).reshape(n, 3)
        return out


class BaryonifyShell(_ShellBase):
    def offsets(self):
        s = self._halo_scalars()
        NSIDE = self.LightconeShell.NSIDE
        T = displacement_table(self.model)
        cut, shift = _model_cut(self.model, s["M"], s["a"])
        off = np.zeros((self.LightconeShell.map.size, 3))
        self.npairs = _lib().bfo_baryonify_shell_offsets(
            NSIDE, s["M"].size, s["vec"], s["lon"], s["lat"], s["radius"], s["D"], s["a"], s["z_in"],
            s["M_in"], cut, shift, *T.args(), off, self.nthreads)
        return off

    def process(self):
        orig = _c(self.LightconeShell.map)
        off = self.offsets()
        new_map = np.zeros(orig.size)
        _lib().bfo_regrid_shell(self.LightconeShell.NSIDE, orig, off, new_map)
        new_sum, old_sum = np.sum(new_map), np.sum(orig)  # HealpixRunner.py:344-346
        assert np.isclose(new_sum, old_sum), \
            "ERROR in pixel regridding, sum(new_map) [%0.14e] != sum(oldmap) [%0.14e]" % (new_sum, old_sum)
        return new_map


class PaintProfilesShell(_ShellBase):
    def process(self):
        assert self.model is not None, "You must provide a model"
        s = self._halo_scalars()
        T = profile_table(self.model, "2D")
        new_map = np.zeros(self.LightconeShell.map.size)
        with extra_axes(self.model, self.HaloLightConeCatalog.cat):
            self.npairs = _lib().bfo_paint_shell(
                self.LightconeShell.NSIDE, s["M"].size, s["vec"], s["radius"], s["D"], s["a"], s["z_in"],
                s["M_in"], *T.args(), new_map, self.nthreads)
        return new_map


# ----------------------------------------------------------------------------
# Grid runners (Runners/Map2DRunner.py)
# ----------------------------------------------------------------------------
class _GridBase:
    def __init__(self, HaloNDCatalog, GriddedMap, epsilon_max, model, use_ellipticity=False,
                 mass_def=None, verbose=True, nthreads=1):
        self.HaloNDCatalog, self.GriddedMap = HaloNDCatalog, GriddedMap
        self.cosmo = HaloNDCatalog.cosmology
        self.model, self.epsilon_max = model, epsilon_max
        self.mass_def = mass_def if mass_def is not None else _ccl().halos.massdef.MassDef(200, "critical")
        self.verbose, self.use_ellipticity, self.nthreads = verbose, use_ellipticity, nthreads
        self.npairs = None
        if use_ellipticity:  # Map2DRunner.py:266-272
            names = HaloNDCatalog.cat.dtype.names
            assert "q_ell" in names, "The 'q_ell' column is missing, but you set use_ellipticity = True"
            if not GriddedMap.is2D:
                assert "c_ell" in names, "The 'c_ell' column is missing, but you set use_ellipticity = True"
            assert "A_ell" in names, "The 'A_ell' column is missing, but you set use_ellipticity = True"

    @staticmethod
    def build_Rmat(A, q):
        """Map2DRunner.py:275-344, 2-D branch, expression by expression (A, q keep the catalogue's float32)."""
        A /= np.linalg.norm(A)
        ref = np.array([1., 0.])
        beta = np.arccos(np.dot(A, ref))
        eta = -np.log(q)
        if eta > 1e-4:
            eta2g = np.tanh(0.5 * eta) / eta
        else:
            etasq = eta * eta
            eta2g = 0.5 + etasq * ((-1 / 24) + etasq * (1 / 240))
        g = eta2g * eta * np.exp(2j * beta)
        g1, g2 = g.real, g.imag
        det = np.sqrt(1 - np.abs(g) ** 2)
        return np.array([[1 + g1, g2], [g2, 1 - g1]]) / det

    def _rmat(self, baryonify):
        """Per-halo shear matrices [n, 4] (row-major), or None.  3-D raises like the reference (:565 / :790)."""
        if not self.use_ellipticity:
            return None
        if not self.GriddedMap.is2D:
            if baryonify:
                raise NotImplementedError("Currently not able to ellipticities with 3D maps.")
            raise ValueError("use_ellipticity is not implemented for 3D maps")
        cat = self.HaloNDCatalog.cat
        out = np.empty((cat.size, 4))
        for j in range(cat.size):
            q_j = cat["q_ell"][j]  # :489-492 / :726-729
            A_j = cat["A_ell"][j]
            A_j = A_j / np.sqrt(np.sum(A_j ** 2))
            assert q_j > 0, "The axis ratio in halo %d is not positive" % j
            out[j] = self.build_Rmat(A_j, q_j).reshape(-1)
        return out

    def _cosmo(self):
        ccl, c = _ccl(), self.cosmo  # Map2DRunner.py:456-460 (no w0)
        return ccl.Cosmology(Omega_c=c["Omega_m"] - c["Omega_b"], Omega_b=c["Omega_b"], h=c["h"],
                             sigma8=c["sigma8"], n_s=c["n_s"], matter_power_spectrum="linear")

    def _positions(self):
        cat, dim = self.HaloNDCatalog.cat, (2 if self.GriddedMap.is2D else 3)
        cols = [np.asarray(cat[k], dtype=np.float64) for k in ("x", "y", "z")[:dim]]
        return _c(np.stack(cols, axis=1)), dim


class BaryonifyGrid(_GridBase):
    def offsets(self):
        cosmo, G, cat = self._cosmo(), self.GriddedMap, self.HaloNDCatalog.cat
        n = cat.size
        a_j = 1 / (1 + self.HaloNDCatalog.redshift)
        Nsize = np.zeros(n, dtype=np.int64)
        M = cat["M"]  # native dtype ('>f4'): the reference takes np.log / get_radius of the float32 scalar
        for j in range(n):  # Map2DRunner.py:484-497
            R_j = self.mass_def.get_radius(cosmo, cat["M"][j], a_j)
            R_q = self.epsilon_max * R_j / a_j
            R_q = np.clip(R_q, 0, np.max(G.bins) / 2)
            Ns = 2 * R_q / G.res
            Nsize[j] = int(Ns // 2) * 2
        a = np.full(n, a_j)
        cut, shift = _model_cut(self.model, M, a)
        pos, dim = self._positions()
        T = displacement_table(self.model)
        off = np.zeros((G.map.size, dim))
        rmat = self._rmat(True)
        self.npairs = _lib().bfo_grid_halos(dim, 0, G.Npix, _c(G.bins), float(G.res), n, pos, Nsize,
                                            np.log(1 / a), _log_native(M), cut, shift,
                                            None if rmat is None else rmat.ctypes.data, *T.args(), off,
                                            self.nthreads)
        return off

    def process(self):
        G = self.GriddedMap
        off = self.offsets()
        orig = _c(G.map).reshape(-1)
        new_map = np.zeros(orig.size)
        _lib().bfo_grid_regrid(2 if G.is2D else 3, G.Npix, off, orig, new_map)
        new_sum, old_sum = np.sum(new_map), np.sum(orig)  # Map2DRunner.py:611-613
        assert np.isclose(new_sum, old_sum), \
            "ERROR in pixel regridding, sum(new_map) [%0.14e] != sum(oldmap) [%0.14e]" % (new_sum, old_sum)
        return new_map.reshape(G.map.shape)


class PaintProfilesGrid(_GridBase):
    def process(self):
        cosmo, G, cat = self._cosmo(), self.GriddedMap, self.HaloNDCatalog.cat
        n = cat.size
        a_j = 1 / (1 + self.HaloNDCatalog.redshift)
        Nsize, cut = np.zeros(n, dtype=np.int64), np.empty(n)
        M = cat["M"]
        for j in range(n):  # Map2DRunner.py:722-733, :804
            R_j = self.mass_def.get_radius(cosmo, cat["M"][j], a_j) / a_j
            Ns = 2 * self.epsilon_max * R_j / G.res
            Ns = int(Ns // 2) * 2
            Nsize[j] = np.clip(Ns, 2, G.bins.size // 2)
            cut[j] = R_j * self.epsilon_max
        a = np.full(n, a_j)
        pos, dim = self._positions()
        T = profile_table(self.model, "2D" if G.is2D else "3D")
        out = np.zeros(G.map.size)
        rmat = self._rmat(False)
        with extra_axes(self.model, cat):
            self.npairs = _lib().bfo_grid_halos(dim, 1, G.Npix, _c(G.bins), float(G.res), n, pos, Nsize,
                                                np.log(1 / a), _log_native(M), cut, np.zeros(n),
                                                None if rmat is None else rmat.ctypes.data, *T.args(), out,
                                                self.nthreads)
        return out.reshape(G.map.shape)


# ----------------------------------------------------------------------------
# Snapshot runner (Runners/SnapshotRunner.py)
# ----------------------------------------------------------------------------
class BaryonifySnapshot:
    def __init__(self, HaloNDCatalog, ParticleSnapshot, epsilon_max, model, mass_def=None, verbose=True,
                 KDTree_kwargs={}, nthreads=1):
        self.HaloNDCatalog, self.ParticleSnapshot = HaloNDCatalog, ParticleSnapshot
        self.epsilon_max, self.cosmo, self.model = epsilon_max, HaloNDCatalog.cosmology, model
        self.mass_def = mass_def if mass_def is not None else _ccl().halos.massdef.MassDef(200, "critical")
        self.verbose, self.nthreads, self.npairs = verbose, nthreads, None

    def offsets(self):
        ccl, c = _ccl(), self.cosmo  # SnapshotRunner.py:195-199
        cosmo = ccl.Cosmology(Omega_c=c["Omega_m"] - c["Omega_b"], Omega_b=c["Omega_b"], h=c["h"],
                              sigma8=c["sigma8"], n_s=c["n_s"], matter_power_spectrum="linear")
        S, cat = self.ParticleSnapshot, self.HaloNDCatalog.cat
        dim = 2 if S.is2D else 3
        n = cat.size
        a_j = 1 / (1 + self.HaloNDCatalog.redshift)
        Rq = np.empty(n)
        M = cat["M"]
        for j in range(n):  # :212-215
            R_j = self.mass_def.get_radius(cosmo, cat["M"][j], a_j)
            Rq[j] = np.clip(self.epsilon_max * R_j / a_j, 0, S.L / 2)
        a = np.full(n, a_j)
        cut, shift = _model_cut(self.model, M, a)
        hpos = _c(np.stack([np.asarray(cat[k], dtype=np.float64) for k in ("x", "y", "z")[:dim]], axis=1))
        part = _c(np.stack([S.cat[k] for k in ("x", "y", "z")[:dim]], axis=1))
        T = displacement_table(self.model)
        off = np.zeros((part.shape[0], dim))
        self.npairs = _lib().bfo_snapshot_offsets(dim, part.shape[0], part, float(S.L), n, hpos, Rq, float(a_j),
                                                  np.log(1 / a), _log_native(M), cut, shift, *T.args(), off,
                                                  self.nthreads)
        return off

    def pair_list(self):
        """Sorted uint64 array of ``particle << 32 | halo`` for every pair of the per-halo ``tree.query_ball_point``
        calls (SnapshotRunner.py:219 / :234), by brute force over all particles (no cell list, no tree)."""
        ccl, c = _ccl(), self.cosmo
        cosmo = ccl.Cosmology(Omega_c=c["Omega_m"] - c["Omega_b"], Omega_b=c["Omega_b"], h=c["h"],
                              sigma8=c["sigma8"], n_s=c["n_s"], matter_power_spectrum="linear")
        S, cat = self.ParticleSnapshot, self.HaloNDCatalog.cat
        dim = 2 if S.is2D else 3
        a_j = 1 / (1 + self.HaloNDCatalog.redshift)
        Rq = np.array([np.clip(self.epsilon_max * self.mass_def.get_radius(cosmo, cat["M"][j], a_j) / a_j, 0, S.L / 2)
                       for j in range(cat.size)], dtype=np.float64)
        hpos = _c(np.stack([np.asarray(cat[k], dtype=np.float64) for k in ("x", "y", "z")[:dim]], axis=1))
        part = _c(np.stack([S.cat[k] for k in ("x", "y", "z")[:dim]], axis=1))
        pairs, cap = np.empty(1, dtype=np.uint64), 0
        for _ in range(2):
            n = _lib().bfo_snapshot_pair_list(dim, part.shape[0], part, float(S.L), cat.size, hpos, Rq, pairs, cap,
                                              self.nthreads)
            if n <= cap:
                break
            pairs, cap = np.empty(n, dtype=np.uint64), n
        return np.sort(pairs[:n])

    def process(self):
        S = self.ParticleSnapshot
        off = self.offsets()
        new_cat = S.cat.copy()
        for m, k in enumerate(("x", "y") + (() if S.is2D else ("z",))):  # :250-260
            out = np.empty(off.shape[0])
            _lib().bfo_snapshot_apply(off.shape[0], _c(S.cat[k]), _c(off[:, m]), float(S.L), out)
            new_cat[k] = out
        return new_cat
