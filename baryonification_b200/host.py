"""Vectorised host-side set-up of the per-halo scalars the kernels consume.

The reference computes these inside its Python loop, one halo at a time, through pyccl
(``Runners/HealpixRunner.py:293-305``, ``Runners/Map2DRunner.py:478-497,716-733``,
``Runners/SnapshotRunner.py:207-215``, ``Profiles/BaryonCorrection.py:364-381``).  Here they are
numpy array expressions with the same operation order and the same dtypes (``HaloNDCatalog`` stores
float32, and the reference takes ``np.log`` / ``get_radius`` of those float32 scalars), handed to the
device as float64 arrays.  Table construction and cosmology stay on the host, as in the reference.
"""
from __future__ import annotations

import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
from scipy import interpolate

from ._ccl import get_ccl

# Chunks of the catalogue are set up by a thread pool: numpy ufuncs and scipy's spline evaluation release the GIL,
# every expression is element-wise over halos, so the result is bit-identical to the serial evaluation.
_PAR_MIN = 100_000
_pool = None


def host_cores():
    """Host cores this process may use: its affinity mask, shared evenly with the other ranks of a one-box job
    (LOCAL_WORLD_SIZE, set by torchrun)."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    try:
        n //= max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))
    except ValueError:
        pass
    return max(1, n)


def _threads():
    env = os.environ.get("BF_HOST_THREADS")
    if env:
        return max(1, int(env))
    return max(1, min(16, host_cores()))


def _interval(xb, x):
    """``clip(searchsorted(xb, x, 'right') - 1, 0, n - 2)`` -- the spline's interval rule.  The runners' breakpoints are
    a linspace (HealpixRunner.py:279), so the interval is guessed from the spacing, moved one step either way against
    the STORED breakpoints and verified; the binary search (7 of the spline's 10 ms per 125 000 halos: random queries
    mispredict every level) only runs for a chunk that holds something the guess cannot prove (NaN, a non-uniform axis)."""
    n = xb.size
    h = (xb[-1] - xb[0]) / (n - 1)
    if n > 2 and h > 0:
        with np.errstate(invalid="ignore"):
            i = ((x - xb[0]) * (1.0 / h)).astype(np.int64)
        np.clip(i, 0, n - 2, out=i)
        i -= (x < xb[i]) & (i > 0)
        i += (x >= xb[i + 1]) & (i < n - 2)
        inside = (xb[i] <= x) & (x < xb[i + 1])
        if np.all(inside | ((i == 0) & (x < xb[0])) | ((i == n - 2) & (x >= xb[n - 1]))):
            return i
    return np.clip(np.searchsorted(xb, x, side="right") - 1, 0, n - 2)


def spline_eval(cs, x):
    """``cs(x)`` of a scipy CubicSpline in numpy, bit-identical to scipy's evaluation (same interval rule, same
    power-accumulation order c3 + c2 s + c1 s^2 + c0 s^3 with s^3 = (s s) s) but free of the GIL, so catalogue chunks
    evaluate in parallel.  Checked against scipy in tests/test_host.py."""
    xb, c = cs.x, cs.c
    i = _interval(xb, x)
    s = x - xb[i]
    res = c[3, i] + c[2, i] * s
    s2 = s * s
    res = res + c[1, i] * s2
    return res + c[0, i] * (s2 * s)


def _pinned_outputs(schema, n):
    """One page-locked float64 slab per output (torch's caching host allocator keeps the blocks between calls), as
    numpy views + the backing tensors.  None when there is no CUDA runtime to pin with."""
    try:
        import torch
        if not torch.cuda.is_available():
            return None
        pins = {k: torch.empty((n,) + tuple(tail), dtype=torch.float64, pin_memory=True) for k, tail in schema.items()}
    except Exception:
        return None
    return pins


def chunked(fn, n, schema=None, runs=1, on_run=None):
    """Evaluate ``fn(slice) -> dict of arrays`` over [0, n) in parallel chunks and join them along axis 0.

    With ``schema`` = {key: trailing shape} (all float64) every chunk writes its slice straight into page-locked
    output arrays -- no serial concatenate, and the upload that follows is one asynchronous copy per array at PCIe
    speed instead of a pageable one.  The backing tensors travel under the key ``"_pinned"``.
    ``runs`` > 1 with ``on_run``: the catalogue is worked through in that many runs of consecutive halos (each run
    spread over all threads) and ``on_run(j0, j1, pinned tensors)`` is called after each, so that the caller can send a
    finished run to the device (upload, table collapse, binning) while the host prepares the next one."""
    global _pool
    nt = _threads()
    if n < _PAR_MIN or nt == 1:
        return fn(slice(0, n))
    if _pool is None:
        _pool = ThreadPoolExecutor(max_workers=nt, thread_name_prefix="bf-host")
    pins = _pinned_outputs(schema, n) if schema else None
    if pins is None:
        edges = np.linspace(0, n, nt + 1).astype(np.int64)
        slices = [slice(int(edges[i]), int(edges[i + 1])) for i in range(nt)]
        parts = list(_pool.map(fn, slices))
        return {k: np.concatenate([p[k] for p in parts], axis=0) for k in parts[0]}
    out = {k: t.numpy() for k, t in pins.items()}

    def run(sl):
        r = fn(sl)
        for k in out:
            out[k][sl] = r[k]

    nruns = max(1, int(runs)) if on_run is not None else 1
    bounds = np.linspace(0, n, nruns + 1).astype(np.int64)
    for r0, r1 in zip(bounds[:-1], bounds[1:]):
        edges = np.linspace(int(r0), int(r1), nt + 1).astype(np.int64)
        list(_pool.map(run, [slice(int(edges[i]), int(edges[i + 1])) for i in range(nt) if edges[i + 1] > edges[i]]))
        if on_run is not None:
            on_run(int(r0), int(r1), pins)
    out["_pinned"] = pins
    return out


def runner_cosmology(cosmo: dict, with_w0: bool):
    """ccl.Cosmology as built at HealpixRunner.py:268-272 (with w0) / Map2DRunner.py:456-459 (without)."""
    ccl = get_ccl()
    kw = dict(Omega_c=cosmo["Omega_m"] - cosmo["Omega_b"], Omega_b=cosmo["Omega_b"], h=cosmo["h"],
              sigma8=cosmo["sigma8"], n_s=cosmo["n_s"], matter_power_spectrum="linear")
    if with_w0:
        kw["w0"] = cosmo["w0"]
    return ccl.Cosmology(**kw)


def radius_of(mass_def, cosmo, M, a):
    """``mass_def.get_radius(cosmo, M_j, a_j)`` for all halos.  M keeps its catalogue dtype so numpy's
    promotion is the one the reference's per-halo call sees.  Falls back to a per-halo loop when the
    mass definition does not broadcast over ``a``."""
    M = np.asarray(M)
    a_arr = np.broadcast_to(np.asarray(a, dtype=np.float64), M.shape)
    try:
        R = np.asarray(mass_def.get_radius(cosmo, M, a if np.ndim(a) == 0 else a_arr))
        if R.shape == M.shape:
            return R.astype(np.float64)
    except Exception:
        pass
    return np.array([mass_def.get_radius(cosmo, M[j], a_arr[j]) for j in range(M.size)], dtype=np.float64)


def model_cut(model, M, a, need_cut=True):
    """(eps_model * R_model, ln R_model if Rdelta_sampling else 0) -- BaryonCorrection.py:370-381."""
    n = np.asarray(M).size
    if not need_cut:
        return np.zeros(n), np.zeros(n)
    R = radius_of(model.mass_def, model.cosmo, M, a) / a
    cut = model.epsilon_max * R
    shift = np.log(R) if getattr(model, "Rdelta_sampling", False) else np.zeros(n)
    return np.asarray(cut, dtype=np.float64), np.asarray(shift, dtype=np.float64)


def property_columns(cat, model):
    """Per-halo values of the model's extra table axes, ``o_j = {key: cat[key][j] for key in p_keys}``
    (HealpixRunner.py:425, Map2DRunner.py:720), as float64 [n, n_keys] -- ``kwargs[k] * empty`` (Tabulate.py:607)
    widens the catalogue scalar (float32 in ND catalogues) to float64.  None when the model has no p_keys."""
    keys = list(vars(model).get("p_keys", []) or []) if model is not None else []
    if not keys:
        return None
    return np.ascontiguousarray(np.stack([np.asarray(cat[k]).astype(np.float64) for k in keys], axis=1))


def log_native(M):
    """np.log(M_j) in the catalogue's dtype (BaryonCorrection.py:369, Tabulate.py:283), widened."""
    M = np.asarray(M)
    if M.dtype.byteorder == ">":
        M = M.astype(M.dtype.newbyteorder("="))
    return np.log(M).astype(np.float64)


def ang2vec_lonlat(lon, lat):
    """healpy.ang2vec(lon, lat, lonlat=True)."""
    theta = np.pi / 2.0 - np.radians(lat)
    phi = np.radians(lon)
    st = np.sin(theta)
    return np.ascontiguousarray(np.stack([st * np.cos(phi), st * np.sin(phi), np.cos(theta)], axis=1))


def shell_scalars(catalog, cosmo_dict, mass_def, epsilon_max, model, baryonify: bool, runs=1, on_run=None):
    """Per-halo inputs of the shell kernels.  ``runs`` / ``on_run``: see :func:`chunked` (large catalogues only)."""
    ccl = get_ccl()
    cosmo = runner_cosmology(cosmo_dict, with_w0=True)
    z_t = np.linspace(0, 30, 1000)  # HealpixRunner.py:279-280
    D_a = interpolate.CubicSpline(z_t, ccl.angular_diameter_distance(cosmo, 1 / (1 + z_t)))
    cat = catalog.cat
    if len(cat) > 0:  # force the back-end's lazy initialisation before threads share the cosmology objects
        mass_def.get_radius(cosmo, float(cat["M"][0]), 1.0)
        if baryonify:
            model.mass_def.get_radius(model.cosmo, float(cat["M"][0]), 1.0)

    def part(sl):
        c = cat[sl]
        M, z = c["M"], c["z"]
        a = 1 / (1 + z)
        R = radius_of(mass_def, cosmo, M, a)
        D = spline_eval(D_a, np.asarray(z, dtype=np.float64))
        out = dict(a=a, D=D, radius=R * epsilon_max / D, z_in=np.log(1 / a), M_in=log_native(M),
                   lon=np.asarray(c["ra"], dtype=np.float64), lat=np.asarray(c["dec"], dtype=np.float64))
        out["vec"] = ang2vec_lonlat(out["lon"], out["lat"])
        out["cut"], out["lnr_shift"] = model_cut(model, M, a, need_cut=baryonify)
        return out

    schema = {k: () for k in ("a", "D", "radius", "z_in", "M_in", "lon", "lat", "cut", "lnr_shift")}
    schema["vec"] = (3,)
    return chunked(part, len(cat), schema, runs, on_run)


def _native(a):
    a = np.asarray(a)
    return a.astype(a.dtype.newbyteorder("=")) if a.dtype.byteorder == ">" else a


def shear_matrices(cat):
    """Per-halo 2x2 matrices of ``build_Rmat`` (Map2DRunner.py:275-344, the galsim-style shear of axis ratio
    ``q_ell`` along ``A_ell``) for all halos at once, [n, 4] row-major.  Same expressions and dtypes as the
    reference's per-halo code (the catalogue columns are float32, so eta = -log(q) and the normalisations are
    float32 operations)."""
    q = _native(cat["q_ell"])
    A = _native(cat["A_ell"])
    assert np.all(q > 0), "The axis ratio in halo %d is not positive" % int(np.argmin(q > 0))
    A = A / np.sqrt(np.sum(A ** 2, axis=1))[:, None]                      # :491 / :728
    A = A / np.sqrt(A[:, 0] * A[:, 0] + A[:, 1] * A[:, 1])[:, None]       # build_Rmat: A /= np.linalg.norm(A)
    beta = np.arccos(A[:, 0].astype(np.float64) * 1.0 + A[:, 1].astype(np.float64) * 0.0)  # np.dot(A, [1, 0])
    eta = -np.log(q)
    etasq = eta * eta
    with np.errstate(divide="ignore", invalid="ignore"):
        eta2g = np.where(eta > 1e-4, np.tanh(0.5 * eta) / eta, 0.5 + etasq * ((-1 / 24) + etasq * (1 / 240)))
    g = (eta2g * eta) * np.exp(2j * beta)
    g1, g2 = g.real, g.imag
    det = np.sqrt(1 - np.abs(g) ** 2)
    return np.ascontiguousarray(np.stack([(1 + g1) / det, g2 / det, g2 / det, (1 - g1) / det], axis=1))


def nd_positions(cat, dim):
    return np.ascontiguousarray(np.stack([np.asarray(cat[k], dtype=np.float64) for k in ("x", "y", "z")[:dim]], axis=1))


def grid_scalars(catalog, gmap, mass_def, epsilon_max, model, baryonify: bool):
    cosmo = runner_cosmology(catalog.cosmology, with_w0=False)
    cat = catalog.cat
    M = cat["M"]
    n = M.size
    a_j = 1 / (1 + catalog.redshift)
    a = np.full(n, a_j)
    R = radius_of(mass_def, cosmo, M, a_j)
    res = gmap.res
    if baryonify:  # Map2DRunner.py:485-497
        R_q = np.clip(epsilon_max * R / a_j, 0, np.max(gmap.bins) / 2)
        Ns = 2 * R_q / res
        Nsize = (Ns // 2).astype(np.int64) * 2
        cut, shift = model_cut(model, M, a_j)
    else:  # Map2DRunner.py:723-733, :804
        Rc = R / a_j
        Ns = 2 * epsilon_max * Rc / res
        Nsize = np.clip((Ns // 2).astype(np.int64) * 2, 2, gmap.bins.size // 2)
        if np.any(Nsize % 2):
            raise ValueError("PaintProfilesGrid: bins.size//2 is odd, the reference's cut-out shapes do not match")
        cut, shift = Rc * epsilon_max, np.zeros(n)
    dim = 2 if gmap.is2D else 3
    return dict(pos=nd_positions(cat, dim), Nsize=Nsize, cut=np.asarray(cut, dtype=np.float64), lnr_shift=shift,
                z_in=np.log(1 / a), M_in=log_native(M), dim=dim, a=a_j)


def grid_rmat(catalog, gmap, use_ellipticity, baryonify):
    """Shear matrices for the 2-D grid runners, or None; 3-D raises like the reference (:565 / :790)."""
    if not use_ellipticity:
        return None
    if not gmap.is2D:
        if baryonify:
            raise NotImplementedError("Currently not able to ellipticities with 3D maps.")
        raise ValueError("use_ellipticity is not implemented for 3D maps")
    return shear_matrices(catalog.cat)


def snapshot_scalars(catalog, snap, mass_def, epsilon_max, model):
    cosmo = runner_cosmology(catalog.cosmology, with_w0=False)
    cat = catalog.cat
    M = cat["M"]
    n = M.size
    a_j = 1 / (1 + catalog.redshift)
    R = radius_of(mass_def, cosmo, M, a_j)
    Rq = np.clip(epsilon_max * R / a_j, 0, snap.L / 2)  # SnapshotRunner.py:213-215
    cut, shift = model_cut(model, M, a_j)
    dim = 2 if snap.is2D else 3
    return dict(pos=nd_positions(cat, dim), Rq=np.asarray(Rq, dtype=np.float64), cut=cut, lnr_shift=shift,
                z_in=np.log(1 / np.full(n, a_j)), M_in=log_native(M), dim=dim, a=a_j)
