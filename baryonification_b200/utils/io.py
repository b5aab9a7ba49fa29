"""Catalog / field containers accepted by the B200 Runners.

These are attribute-compatible with the reference's value types
(``BaryonForge/utils/io.py``: ``HaloLightConeCatalog`` :9-140, ``HaloNDCatalog``
:143-287, ``LightconeShell`` :290-375, ``GriddedMap`` :378-487,
``ParticleSnapshot`` :490-670), so a reference object and one of these are
interchangeable as Runner inputs (the Runners only duck-type on ``.cat``,
``.cosmology``, ``.map``, ``.NSIDE``, ``.bins`` ... ).  Differences:

* no dependency on healpy (``NSIDE`` comes from integer arithmetic; reading a
  FITS path needs healpy and raises ImportError without it);
* ``GriddedMap.grid`` / ``.inds`` are built lazily -- the GPU path never needs
  them and at 1024^3 they would cost 34 GB of host memory.
"""
from __future__ import annotations

import math
import warnings

import numpy as np

__all__ = ["HaloLightConeCatalog", "HaloNDCatalog", "LightconeShell", "GriddedMap", "ParticleSnapshot"]

_REQUIRED = ("Omega_m", "sigma8", "h", "Omega_b", "n_s", "w0")


def _check_cosmo(cosmo):
    if cosmo is None or any(k not in cosmo.keys() for k in _REQUIRED):
        raise ValueError("Not all cosmology parameters provided. I need Omega_m, sigma8, h, sigma8, Omega_b, n_s, w0")
    return cosmo


class _CatalogBase:
    _core = ()

    @property
    def data(self):
        return self.cat

    @property
    def cosmology(self):
        return self.cosmo

    def _extras(self, key):
        return {k: self.cat[k][key] for k in self.cat.dtype.names if k not in self._core}

    def __len__(self):
        return self.cat.size


class HaloLightConeCatalog(_CatalogBase):
    """Halos on the lightcone: float64 fields ``M, z, ra, dec`` (+ extras), degrees."""

    _core = ("ra", "dec", "M", "z")

    def __init__(self, ra, dec, M, z, cosmo, **arrays):
        ra = np.atleast_1d(np.asarray(ra))
        dec = np.atleast_1d(np.asarray(dec))
        fields = [("M", "f8"), ("z", "f8"), ("ra", "f8"), ("dec", "f8")] + [(k, "f8") for k in arrays]
        cat = np.zeros(ra.size, dtype=fields)
        if np.any(np.abs(dec) == 90):
            warnings.warn("Some halos found with declination exactly at the poles. Offsetting these by 4e-5 arcsec")
            dec = np.clip(dec.astype(np.float64), -90 + 1e-8, 90 - 1e-8)
        cat["ra"], cat["dec"], cat["z"], cat["M"] = ra, dec, z, M
        for k, v in arrays.items():
            cat[k] = v
        self.cat = cat
        self.cosmo = _check_cosmo(cosmo)

    def __getitem__(self, key):
        c = self.cat
        return HaloLightConeCatalog(ra=c["ra"][key], dec=c["dec"][key], M=c["M"][key], z=c["z"][key],
                                    cosmo=self.cosmo, **self._extras(key))

    def __str__(self):
        c = self.cat
        return (f"HaloLightConeCatalog with {c.size} Halos at {c['z'].min()} < z < {c['z'].max()}.\n"
                f"Minimum log10(Mass) = {np.log10(c['M'].min())}\nMaximum log10(Mass) = {np.log10(c['M'].max())}\n"
                f"Cosmology set to {self.cosmo}.")


class HaloNDCatalog(_CatalogBase):
    """Halos in a periodic box at one redshift.

    Fields ``M, x, y, z`` are stored as **big-endian float32** (``'>f'``), exactly like the
    reference (io.py:204), because that rounding is visible in the results.
    """

    _core = ("x", "y", "z", "M")

    def __init__(self, x, y, M, redshift, cosmo, z=None, **arrays):
        x = np.atleast_1d(np.asarray(x))
        fields = [("M", ">f"), ("x", ">f"), ("y", ">f"), ("z", ">f")]
        for k, v in arrays.items():
            v = np.asarray(v)
            fields.append((k, ">f", v.shape[1:]) if v.ndim > 1 else (k, ">f"))
        cat = np.zeros(x.size, dtype=fields)
        cat["x"], cat["y"], cat["M"] = x, y, M
        cat["z"] = 0 if z is None else z
        for k, v in arrays.items():
            cat[k] = v
        self.cat = cat
        self.redshift = redshift
        self.cosmo = _check_cosmo(cosmo)

    def __getitem__(self, key):
        c = self.cat
        return HaloNDCatalog(x=c["x"][key], y=c["y"][key], z=c["z"][key], M=c["M"][key],
                             redshift=self.redshift, cosmo=self.cosmo, **self._extras(key))

    def __str__(self):
        c = self.cat
        return (f"HaloNDCatalog with {c.size} Halos at z = {self.redshift}.\n"
                f"Minimum log10(Mass) = {np.log10(c['M'].min())}\nMaximum log10(Mass) = {np.log10(c['M'].max())}\n"
                f"Cosmology set to {self.cosmo}.")

    def __repr__(self):
        return f"HaloNDCatalog(cat = {self.cat!r}, \nredshift = {self.redshift!r}, \ncosmo = {self.cosmo})"


def npix2nside(npix: int) -> int:
    nside = math.isqrt(int(npix) // 12)
    if 12 * nside * nside != int(npix):
        raise ValueError("Wrong pixel number (it is not 12*nside**2)")
    return nside


class LightconeShell:
    """A full-sky HEALPix map in RING ordering plus cosmology."""

    def __init__(self, map=None, path=None, cosmo=None):
        if path is None and map is None:
            raise ValueError("Need to provide either path to map, or provide map values in healpix ring configuration")
        if isinstance(path, str):
            try:
                import healpy as hp
            except Exception as e:  # pragma: no cover - healpy is not in the build image
                raise ImportError("reading a FITS map needs healpy; pass map=<ndarray> instead") from e
            self.map = hp.read_map(path)
        elif isinstance(map, np.ndarray):
            self.map = map
        self.NSIDE = npix2nside(self.map.size)
        self.cosmo = _check_cosmo(cosmo)

    @property
    def data(self):
        return self.map

    @property
    def cosmology(self):
        return self.cosmo


class GriddedMap:
    """A square / cubic periodic grid.  ``bins`` are the pixel-centre coordinates along one axis."""

    def __init__(self, map=None, redshift=None, bins=None, cosmo=None):
        self.map = map
        self.redshift = redshift
        self.Npix = self.map.shape[0]
        self.res = bins[1] - bins[0]
        self.bins = bins
        self.is2D = len(self.map.shape) == 2
        assert all(s == self.map.shape[0] for s in self.map.shape), "maps have to be square / cubic"
        self._grid = None
        self._inds = None
        self.cosmo = _check_cosmo(cosmo)

    @property
    def grid(self):
        if self._grid is None:
            self._grid = np.meshgrid(*([self.bins] * (2 if self.is2D else 3)), indexing="xy")
        return self._grid

    @property
    def inds(self):
        if self._inds is None:
            self._inds = np.arange(self.map.size).reshape(self.map.shape)
        return self._inds

    @property
    def data(self):
        return self.map

    @property
    def cosmology(self):
        return self.cosmo


class ParticleSnapshot:
    """Particles in a periodic box of side ``L``: float64 fields ``M, x, y, z``."""

    def __init__(self, x=None, y=None, z=None, M=None, L=None, redshift=None, cosmo=None):
        cat = np.zeros(len(x), dtype=[("M", "f8"), ("x", "f8"), ("y", "f8"), ("z", "f8")])
        cat["x"], cat["y"], cat["M"] = x, y, M
        cat["z"] = 0 if z is None else z
        self.L = L
        self.cat = cat
        self.redshift = redshift
        self.is2D = z is None
        self.cosmo = _check_cosmo(cosmo)

    @property
    def data(self):
        return self.cat

    @property
    def cosmology(self):
        return self.cosmo

    def make_map(self, N_grid, device=None):
        """Mass map on an ``N_grid``-per-side mesh: ``np.histogramdd(coords, bins=(linspace(0, L, N_grid + 1),) * D,
        weights=M)[0]`` (utils/io.py:657-670) evaluated on the GPU by ``bf_make_map`` -- the same bins (the numpy edges
        are uploaded and searched exactly) and the same per-cell summation order (particle order), hence the same bits.
        ``device="host"`` runs numpy itself (no GPU needed: what the reference does)."""
        assert np.isnan(self.cat["M"]).sum() == 0, "If you want to make a map, provide a value for the particle mass"
        edges = np.linspace(0, self.L, N_grid + 1)
        names = ("x", "y") if self.is2D else ("x", "y", "z")
        if device == "host":
            coords = np.stack([self.cat[k] for k in names], axis=1)
            return np.histogramdd(coords, bins=(edges,) * len(names), weights=self.cat["M"])[0]
        import ctypes as C
        import torch
        from .. import _device as dv, _lib
        from ..Runners.snapshot import BaryonifySnapshot
        lib = _lib.load(require_cuda=True)
        dev = dv.device(device)
        dim, n = len(names), len(self.cat)
        with torch.cuda.device(dev):
            grid = dv.empty((int(N_grid),) * dim, dev=dev)
            ws = torch.empty(max(lib.bf_make_map_ws_bytes(n), 256), dtype=torch.uint8, device=dev)
            e_d = dv.to_dev(edges, dev=dev)
            if BaryonifySnapshot._is_reference_layout(self.cat):
                rec = dv.AsyncUpload(self.cat.view(np.float64).reshape(-1, 4), dev).result()
                args = (None, None, None, None, dv.ptr(rec))
            else:
                cols = [dv.to_dev(self.cat[k], dev=dev, pinned=True) for k in names]
                cols += [None] * (3 - dim) + [dv.to_dev(self.cat["M"], dev=dev, pinned=True)]
                args = tuple(dv.ptr(c) for c in cols) + (None,)
            _lib.call("bf_make_map", dim, n, *args, dv.ptr(e_d), int(N_grid), dv.ptr(grid), dv.ptr(ws), ws.numel(),
                      dv.stream_ptr())
            return dv.to_host(grid, True)
