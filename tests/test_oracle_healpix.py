"""CPU: pins the HEALPix RING restatement (oracle/healpix_oracle.c) with hand-derivable known answers and
brute-force properties.  healpy itself is not installable here ("parity unpinned" against it); when
`import healpy` does succeed these same tests run against the real library through oracle.ref_shim."""
import numpy as np
import pytest

from oracle import healpix as ohp


def test_nside1_known_answers():
    # 12 base pixels: rings at z = 2/3, 0, -2/3; phi = pi/4 + k pi/2 (caps), k pi/2 (equator)
    x, y, z = ohp.pix2vec(1, np.arange(12))
    assert np.allclose(z, [2 / 3] * 4 + [0] * 4 + [-2 / 3] * 4, atol=1e-15)
    phi = np.mod(np.arctan2(y, x), 2 * np.pi)
    want = np.concatenate([np.pi / 4 + np.arange(4) * np.pi / 2, np.arange(4) * np.pi / 2,
                           np.pi / 4 + np.arange(4) * np.pi / 2])
    assert np.allclose(phi, want, atol=1e-15)
    assert np.allclose(x * x + y * y + z * z, 1.0, atol=1e-15)


@pytest.mark.parametrize("nside", [1, 2, 3, 4, 16, 64])
def test_ring_starts_and_latitudes(nside):
    npix, ncap = 12 * nside * nside, 2 * nside * (nside - 1)
    th, ph = ohp.pix2ang(nside, np.arange(npix))
    z = np.cos(th)
    for i in range(1, 4 * nside):  # ring i: first pixel, number of pixels, z
        if i < nside:
            start, n, zi = 2 * i * (i - 1), 4 * i, 1 - i * i / (3.0 * nside * nside)
        elif i <= 3 * nside:
            start, n, zi = ncap + 4 * nside * (i - nside), 4 * nside, (2 * nside - i) * 2.0 / (3.0 * nside)
        else:
            k = 4 * nside - i
            start, n, zi = npix - 2 * k * (k + 1), 4 * k, -1 + k * k / (3.0 * nside * nside)
        assert np.allclose(z[start:start + n], zi, atol=2e-15)
        assert np.all(np.diff(ph[start:start + n]) > 0)
        shifted = i < nside or i > 3 * nside or ((i - nside) % 2 == 0)
        assert np.isclose(ph[start], (np.pi / n) if shifted else 0.0, atol=1e-15)


@pytest.mark.parametrize("nside", [1, 2, 8, 32, 1024, 4096])
def test_pix_ang_roundtrip(nside):
    npix = 12 * nside * nside
    rng = np.random.default_rng(nside)
    pix = np.arange(npix) if npix <= 20000 else np.unique(np.concatenate(
        [rng.integers(0, npix, 20000), np.arange(500), npix - 1 - np.arange(500)]))
    th, ph = ohp.pix2ang(nside, pix)
    assert np.array_equal(ohp.ang2pix(nside, th, ph), pix)
    x, y, z = ohp.pix2vec(nside, pix)
    lon, lat = ohp.vec2ang(np.stack([x, y, z], 1), lonlat=True)
    assert np.array_equal(ohp.ang2pix(nside, lon, lat, lonlat=True), pix)


def test_npix2nside_and_resol():
    assert ohp.npix2nside(12 * 4096 ** 2) == 4096
    with pytest.raises(ValueError):
        ohp.npix2nside(13)
    assert np.isclose(ohp.nside2resol(64), np.sqrt(4 * np.pi / (12 * 64 ** 2)), rtol=1e-15)


@pytest.mark.parametrize("nside", [1, 2, 4, 16, 64])
def test_query_disc_equals_brute_force(nside):
    """query_disc(inclusive=False) == {p : angle(centre_p, v) < r}, away from 1-ulp edge cases."""
    rng = np.random.default_rng(17 + nside)
    npix = 12 * nside * nside
    vecs = np.stack(ohp.pix2vec(nside, np.arange(npix)), 1)
    n = 200
    c = rng.normal(size=(n, 3))
    c /= np.linalg.norm(c, axis=1)[:, None]
    c[0], c[1] = [0, 0, 1], [0, 0, -1]
    c[2] = [1, 0, 0]
    c[3] = np.array([np.cos(6.2831), np.sin(6.2831), 0.0])  # just below phi = 2 pi: wrap-around
    rad = np.where(rng.random(n) < 0.7, rng.uniform(0, 0.4, n), rng.uniform(0, 3.3, n))
    for j in range(n):
        got = ohp.query_disc(nside, c[j], rad[j])
        ang = np.arccos(np.clip(vecs @ c[j], -1, 1))
        eps = 1e-9
        must = np.flatnonzero(ang < rad[j] - eps)
        may = np.flatnonzero(ang < rad[j] + eps)
        assert np.all(np.diff(got) > 0), "ascending, unique"
        assert np.isin(must, got).all() and np.isin(got, may).all(), j
    assert ohp.query_disc(nside, c[5], 4.0).size == npix  # radius >= pi: everything
    assert ohp.query_disc(nside, c[5], 0.0).size == 0


@pytest.mark.parametrize("nside", [1, 4, 32, 512])
def test_interp_weights_properties(nside):
    rng = np.random.default_rng(99 + nside)
    n = 5000
    lon = rng.uniform(0, 360, n)
    lat = np.degrees(np.arcsin(rng.uniform(-1, 1, n)))
    lat[:40] = np.concatenate([np.linspace(89.0, 89.99999, 20), -np.linspace(89.0, 89.99999, 20)])  # polar caps
    pix, w = ohp.get_interp_weights(nside, lon, lat, lonlat=True)
    assert pix.shape == (4, n) and w.shape == (4, n)
    assert np.allclose(w.sum(0), 1.0, atol=1e-12)  # what the mass-conservation assert relies on
    assert (w > -1e-12).all() and (w < 1 + 1e-12).all()
    assert (pix >= 0).all() and (pix < 12 * nside * nside).all()
    # at a pixel centre the weight collapses onto that pixel
    p0 = rng.integers(0, 12 * nside * nside, 300)
    l0, b0 = ohp.pix2ang(nside, p0, lonlat=True)
    pp, ww = ohp.get_interp_weights(nside, l0, b0, lonlat=True)
    assert np.allclose(np.where(pp == p0[None, :], ww, 0).sum(0), 1.0, atol=1e-6)
    # bilinear interpolation reproduces a function linear in z exactly between two rings (belt)
    if nside >= 4:
        th = np.pi / 2 - np.radians(lat)
        sel = np.abs(np.cos(th)) < 0.6
        thp, _ = ohp.pix2ang(nside, pix[:, sel].ravel())
        est = (w[:, sel] * thp.reshape(4, -1)).sum(0)
        assert np.allclose(est, th[sel], atol=1e-10)


def test_scalar_call_shapes():
    # the call shapes the reference uses (HealpixRunner.py:303,306,310)
    v = ohp.ang2vec(10.0, 20.0, lonlat=True)
    assert v.shape == (3,)
    p, w = ohp.get_interp_weights(16, 10.0, 20.0, lonlat=True)
    assert p.shape == (4,) and w.shape == (4,)
    assert ohp.query_disc(16, v, 0.2, inclusive=False, nest=False).dtype == np.int64
