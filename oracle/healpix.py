"""oracle/healpix.py -- TEST INFRASTRUCTURE ONLY.

ctypes front-end of ``oracle/healpix_oracle.c`` shaped like the slice of the
``healpy`` API the reference's Shell runners call (HealpixRunner.py:303-337,
429-433; io.py:350; Pixel.py:479).  PARITY UNPINNED against real healpy (it is
not installable here); see the header of healpix_oracle.c.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_i64 = C.c_int64
_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")


def build(force: bool = False) -> str:
    """Compile liboracle (gcc).  Building the checker is not using it."""
    so = os.path.join(_HERE, "libbf_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("healpix_oracle.c", "runners_oracle.c")]
    if force or (not os.path.exists(so)) or any(
            os.path.getmtime(s) > os.path.getmtime(so) for s in srcs if os.path.exists(s)):
        subprocess.check_call(["make", "-s", "-C", _HERE, "libbf_oracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.hpxo_npix2nside.restype = _i64
        L.hpxo_npix2nside.argtypes = [_i64]
        L.hpxo_nside2resol.restype = C.c_double
        L.hpxo_nside2resol.argtypes = [_i64]
        L.hpxo_ang2vec_lonlat.argtypes = [_i64, _dp, _dp, _dp]
        L.hpxo_vec2ang_lonlat.argtypes = [_i64, _dp, _dp, _dp]
        L.hpxo_pix2vec_ring.argtypes = [_i64, _i64, _ip, _dp]
        L.hpxo_pix2ang_ring.argtypes = [_i64, _i64, _ip, _dp, _dp]
        L.hpxo_ang2pix_ring.argtypes = [_i64, _i64, _dp, _dp, _ip]
        L.hpxo_query_disc_ranges.restype = _i64
        L.hpxo_query_disc_ranges.argtypes = [_i64, _dp, C.c_double, _i64, _ip]
        L.hpxo_get_interp_weights_lonlat.argtypes = [_i64, _i64, _dp, _dp, _ip, _dp]
        L.hpxo_get_interp_weights.argtypes = [_i64, _i64, _dp, _dp, _ip, _dp]
        _LIB = L
    return _LIB


def _f(a):
    return np.ascontiguousarray(np.atleast_1d(a), dtype=np.float64)


def nside2npix(nside):
    return 12 * int(nside) * int(nside)


def npix2nside(npix):
    ns = lib().hpxo_npix2nside(int(npix))
    if ns < 0:
        raise ValueError("Wrong pixel number (it is not 12*nside**2)")
    return int(ns)


def nside2resol(nside, arcmin=False):
    r = lib().hpxo_nside2resol(int(nside))
    return np.rad2deg(r) * 60.0 if arcmin else r


def ang2vec(theta, phi, lonlat=False):
    scalar = np.ndim(theta) == 0 and np.ndim(phi) == 0
    if not lonlat:  # convert (theta,phi) -> (lon,lat) would lose bits; do it directly
        t, p = _f(theta), _f(phi)
        st = np.sin(t)
        out = np.stack([st * np.cos(p), st * np.sin(p), np.cos(t)], axis=1)
    else:
        lon, lat = np.broadcast_arrays(_f(theta), _f(phi))
        lon, lat = _f(lon), _f(lat)
        out = np.empty((lon.size, 3))
        lib().hpxo_ang2vec_lonlat(lon.size, lon, lat, out)
    return out[0] if scalar else out


def vec2ang(vectors, lonlat=False):
    v = np.ascontiguousarray(np.asarray(vectors, dtype=np.float64).reshape(-1, 3))
    if lonlat:
        lon, lat = np.empty(v.shape[0]), np.empty(v.shape[0])
        lib().hpxo_vec2ang_lonlat(v.shape[0], v, lon, lat)
        return lon, lat
    dn = np.sqrt(np.sum(np.square(v), axis=1))
    theta = np.arccos(v[:, 2] / dn)
    phi = np.arctan2(v[:, 1], v[:, 0])
    phi[phi < 0] += 2 * np.pi
    return theta, phi


def pix2vec(nside, ipix, nest=False):
    assert not nest, "oracle restates the RING scheme only"
    scalar = np.ndim(ipix) == 0
    p = np.ascontiguousarray(np.atleast_1d(ipix), dtype=np.int64)
    out = np.empty((p.size, 3))
    lib().hpxo_pix2vec_ring(int(nside), p.size, p, out)
    if scalar:
        return out[0, 0], out[0, 1], out[0, 2]
    return out[:, 0].copy(), out[:, 1].copy(), out[:, 2].copy()


def pix2ang(nside, ipix, nest=False, lonlat=False):
    assert not nest
    p = np.ascontiguousarray(np.atleast_1d(ipix), dtype=np.int64)
    th, ph = np.empty(p.size), np.empty(p.size)
    lib().hpxo_pix2ang_ring(int(nside), p.size, p, th, ph)
    if lonlat:
        return np.degrees(ph), 90.0 - np.degrees(th)
    return th, ph


def ang2pix(nside, theta, phi, nest=False, lonlat=False):
    assert not nest
    scalar = np.ndim(theta) == 0 and np.ndim(phi) == 0
    t, p = np.broadcast_arrays(_f(theta), _f(phi))
    t, p = _f(t), _f(p)
    if lonlat:
        t, p = np.pi / 2.0 - np.radians(p), np.radians(t)
        t, p = _f(t), _f(p)
    out = np.empty(t.size, dtype=np.int64)
    lib().hpxo_ang2pix_ring(int(nside), t.size, t, p, out)
    return int(out[0]) if scalar else out


def query_disc_ranges(nside, vec, radius):
    v = _f(vec)
    cap = 8 * int(nside) + 8
    buf = np.empty(2 * cap, dtype=np.int64)
    n = lib().hpxo_query_disc_ranges(int(nside), v, float(radius), cap, buf)
    if n < 0:
        cap = -n
        buf = np.empty(2 * cap, dtype=np.int64)
        n = lib().hpxo_query_disc_ranges(int(nside), v, float(radius), cap, buf)
    return buf[: 2 * n].reshape(n, 2)


def query_disc(nside, vec, radius, inclusive=False, fact=4, nest=False, buff=None):
    assert not inclusive and not nest, "oracle restates query_disc(inclusive=False, nest=False) only"
    rg = query_disc_ranges(nside, vec, radius)
    if rg.shape[0] == 0:
        return np.empty(0, dtype=np.int64)
    return np.concatenate([np.arange(a, b, dtype=np.int64) for a, b in rg])


def get_interp_weights(nside, theta, phi=None, nest=False, lonlat=False):
    assert not nest
    t, p = np.broadcast_arrays(_f(theta), _f(phi))
    t, p = _f(t), _f(p)
    pix = np.empty((t.size, 4), dtype=np.int64)
    wgt = np.empty((t.size, 4))
    if lonlat:
        lib().hpxo_get_interp_weights_lonlat(int(nside), t.size, t, p, pix, wgt)
    else:
        lib().hpxo_get_interp_weights(int(nside), t.size, t, p, pix, wgt)
    if np.ndim(theta) == 0 and np.ndim(phi) == 0:
        return pix[0].copy(), wgt[0].copy()
    return np.ascontiguousarray(pix.T), np.ascontiguousarray(wgt.T)


def as_module(name="healpy"):
    import types
    m = types.ModuleType(name)
    for k in ("nside2npix", "npix2nside", "nside2resol", "ang2vec", "vec2ang", "pix2vec",
              "pix2ang", "ang2pix", "query_disc", "get_interp_weights"):
        setattr(m, k, globals()[k])
    m.__version__ = "0+oracle-restated"

    def read_map(*a, **k):
        raise NotImplementedError("FITS I/O is out of scope for the oracle")
    m.read_map = read_map
    return m
