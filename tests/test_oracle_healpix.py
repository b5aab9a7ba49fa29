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


# ---------------------------------------------------------------------------------------------------------
# Known answers PUBLISHED by healpy itself: the examples printed in the docstrings of healpy.pixelfunc
# (pix2ang, ang2pix, pix2vec, get_interp_weights, nside2resol, npix2nside) -- outputs of the real library, so they
# pin the restatement against healpy and not only against its author's reading of the algorithm.
# ---------------------------------------------------------------------------------------------------------
def test_healpy_docstring_known_answers():
    # >>> hp.pix2ang(16, 1440)                     -> (1.5291175943723188, 0.0)
    th, ph = ohp.pix2ang(16, np.array([1440]))
    assert th[0] == 1.5291175943723188 and ph[0] == 0.0
    # >>> hp.pix2ang(16, [1440, 427, 1520, 0, 3068]) -> theta (1.52911759, 0.78550497, 1.57079633, 0.05103658, 3.09055608)
    th, ph = ohp.pix2ang(16, np.array([1440, 427, 1520, 0, 3068]))
    assert np.allclose(th, [1.52911759, 0.78550497, 1.57079633, 0.05103658, 3.09055608], rtol=0, atol=5e-9)
    assert np.allclose(ph[:4], [0.0, 0.78539816, 1.61988371, 0.78539816], rtol=0, atol=5e-9)
    # >>> hp.pix2ang([1, 2, 4, 8], 11) -> (2.30052398, 0.84106867, 0.41113786, 0.2044802), (5.49778714, 5.89048623 x 3)
    got = [ohp.pix2ang(ns, np.array([11])) for ns in (1, 2, 4, 8)]
    assert np.allclose([g[0][0] for g in got], [2.30052398, 0.84106867, 0.41113786, 0.2044802], rtol=0, atol=5e-9)
    assert np.allclose([g[1][0] for g in got], [5.49778714, 5.89048623, 5.89048623, 5.89048623], rtol=0, atol=5e-9)
    # >>> hp.ang2pix(16, np.pi/2, 0) -> 1440 ; the array example -> [1440, 427, 1520, 0, 3068]
    th = np.array([np.pi / 2, np.pi / 4, np.pi / 2, 0, np.pi])
    ph = np.array([0.0, np.pi / 4, np.pi / 2 + 1e-15, 0, 0])
    assert np.array_equal(ohp.ang2pix(16, th, ph), [1440, 427, 1520, 0, 3068])
    # >>> hp.pix2vec(16, 1504) -> (0.99879545620517241, 0.049067674327418015, 0.0)
    x, y, z = ohp.pix2vec(16, np.array([1504]))
    assert (x[0], y[0], z[0]) == (0.99879545620517241, 0.049067674327418015, 0.0)
    # >>> hp.pix2vec(16, [1440, 427]) -> ([0.99913157, 0.5000534], [0., 0.5000534], [0.04166667, 0.70703125])
    x, y, z = ohp.pix2vec(16, np.array([1440, 427]))
    assert np.allclose(x, [0.99913157, 0.5000534], atol=5e-9) and np.allclose(y, [0.0, 0.5000534], atol=5e-9)
    assert np.allclose(z, [0.04166667, 0.70703125], atol=5e-9)
    # >>> hp.nside2resol(128, arcmin=True) -> 27.483891294539248 ; hp.nside2resol(256) -> 0.0039973699529159707
    assert np.isclose(ohp.nside2resol(128, arcmin=True), 27.483891294539248, rtol=1e-15)
    assert np.isclose(ohp.nside2resol(256), 0.0039973699529159707, rtol=1e-15)
    # >>> hp.npix2nside(768) -> 8 ; hp.nside2npix(8) -> 768
    assert ohp.npix2nside(768) == 8 and ohp.nside2npix(8) == 768


def test_healpy_docstring_interp_weights():
    """The three examples of healpy.get_interp_weights' docstring: a pixel centre, the north pole (cap pixels
    (pix + 2) & 3, weights 1/4) with the phi = 0 wrap (i1 = -1 -> 3), and the equator with the wrap on ring 3."""
    # >>> hp.get_interp_weights(1, 0)       (one argument: the position of pixel 0)
    # (array([0, 1, 4, 5]), array([ 1.,  0.,  0.,  0.]))
    th, ph = ohp.pix2ang(1, np.array([0]))
    p, w = ohp.get_interp_weights(1, th, ph)
    assert np.array_equal(p[:, 0], [0, 1, 4, 5]) and np.allclose(w[:, 0], [1, 0, 0, 0], atol=1e-15)
    # >>> hp.get_interp_weights(1, 0, 0)
    # (array([1, 2, 3, 0]), array([ 0.25,  0.25,  0.25,  0.25]))
    p, w = ohp.get_interp_weights(1, 0.0, 0.0)
    assert np.array_equal(p, [1, 2, 3, 0]) and np.array_equal(w, [0.25, 0.25, 0.25, 0.25])
    # >>> hp.get_interp_weights(1, [0, np.pi/2], 0)
    # (array([[ 1,  4], [ 2,  5], [ 3, 11], [ 0,  8]]), array([[ 0.25,  1.], [ 0.25,  0.], [ 0.25,  0.], [ 0.25,  0.]]))
    p, w = ohp.get_interp_weights(1, np.array([0.0, np.pi / 2]), np.array([0.0, 0.0]))
    assert np.array_equal(p, [[1, 4], [2, 5], [3, 11], [0, 8]])
    assert np.allclose(w, [[0.25, 1.0], [0.25, 0.0], [0.25, 0.0], [0.25, 0.0]], atol=1e-15)


def test_interp_weights_hand_derived_nside2():
    """Worked by hand from the RING layout of NSIDE 2 (rings of 4, 8, 8, 8, 8, 8, 4 pixels starting at 0, 4, 12, 20,
    28, 36, 44; z = 11/12, 2/3, 1/3, 0, ...; rings 1, 2, 4, 6, 7 start at phi = dphi / 2)."""
    t1, t2 = np.arccos(11 / 12), np.arccos(2 / 3)
    # A: theta = 0.5 (between rings 1 and 2), phi = 0.1 (left of the first pixel centre on both rings: wraps)
    theta, phi = 0.5, 0.1
    a1 = phi / (np.pi / 2) + 0.5   # ring 1: distance from the centre of pixel 3 (at -pi/4), in pixels
    a2 = phi / (np.pi / 4) + 0.5   # ring 2: distance from the centre of pixel 11 (at -pi/8)
    wt = (theta - t1) / (t2 - t1)
    p, w = ohp.get_interp_weights(2, theta, phi)
    assert np.array_equal(p, [3, 0, 11, 4])
    assert np.allclose(w, [(1 - a1) * (1 - wt), a1 * (1 - wt), (1 - a2) * wt, a2 * wt], rtol=0, atol=1e-14)
    # B: theta = 3.0 (south of the last ring, theta_7 = pi - t1), phi = 4.0: pixels 46, 47 of ring 7 + the cap pair
    theta, phi = 3.0, 4.0
    b = phi / (np.pi / 2) - 0.5 - 2  # distance from the centre of pixel 46 (third of the ring, at 5 pi / 4)
    wt = (theta - (np.pi - t1)) / t1
    p, w = ohp.get_interp_weights(2, theta, phi)
    assert np.array_equal(p, [46, 47, 44, 45])
    assert np.allclose(w, [(1 - b) * (1 - wt) + wt / 4, b * (1 - wt) + wt / 4, wt / 4, wt / 4], rtol=0, atol=1e-14)
    # C: belt, rings 3 (unshifted, start 12) and 4 (shifted, start 20): theta = 1.4, phi = 6.2 (wraps on ring 4 only)
    theta, phi = 1.4, 6.2
    t3, t4 = np.arccos(1 / 3), np.pi / 2
    c3 = phi / (np.pi / 4) - 7        # ring 3: between pixel 7 (at 7 pi / 4) and pixel 0 (at 2 pi)
    c4 = phi / (np.pi / 4) - 0.5 - 7  # ring 4: between pixel 7 (at 15 pi / 8) and pixel 0 (at pi / 8 + 2 pi)
    wt = (theta - t3) / (t4 - t3)
    p, w = ohp.get_interp_weights(2, theta, phi)
    assert np.array_equal(p, [12 + 7, 12 + 0, 20 + 7, 20 + 0])
    assert np.allclose(w, [(1 - c3) * (1 - wt), c3 * (1 - wt), (1 - c4) * wt, c4 * wt], rtol=0, atol=1e-14)


@pytest.mark.parametrize("nside", [1024, 4096])
def test_query_disc_equals_brute_force_large_nside(nside):
    """At the benchmark's NSIDE: the ring walk against a dot-product test over every pixel of the rings the disc can
    reach (pix2vec shares no code with the walk).  Polar caps, a disc that just touches a pole, the phi = 0 wrap."""
    rng = np.random.default_rng(900 + nside)
    pixrad = np.sqrt(4 * np.pi / (12 * nside * nside))
    cases = [([0, 0, 1], 100 * pixrad), ([0, 0, -1], 40 * pixrad), ([1e-9, 0, 1], 3 * pixrad),
             ([1.0, -1e-12, 0.3], 60 * pixrad)]
    for k in range(8):
        r = pixrad * rng.choice([0.4, 2.5, 30.0, 150.0]) * rng.uniform(0.8, 1.2)
        d = r * (0.7 if k % 2 else 1.0)  # swallow / touch a pole
        sgn, ph = (1.0 if k % 4 < 2 else -1.0), rng.uniform(0, 2 * np.pi)
        cases.append(([np.sin(d) * np.cos(ph), np.sin(d) * np.sin(ph), sgn * np.cos(d)], r))
    for k in range(8):
        v = rng.normal(size=3)
        cases.append((v, pixrad * rng.choice([0.4, 1.0, 7.0, 150.0]) * rng.uniform(0.8, 1.2)))
    for v, r in cases:
        v = np.asarray(v, dtype=np.float64)
        v = v / np.linalg.norm(v)
        got = ohp.query_disc(nside, v, r)
        assert np.all(np.diff(got) > 0)
        th0 = np.arccos(np.clip(v[2], -1, 1))
        t_lo, t_hi = max(th0 - r - 3 * pixrad, 0.0), min(th0 + r + 3 * pixrad, np.pi)
        p_lo = int(np.atleast_1d(ohp.ang2pix(nside, np.array([t_lo]), np.array([0.0])))[0])
        p_hi = int(np.atleast_1d(ohp.ang2pix(nside, np.array([t_hi]), np.array([0.0])))[0])
        cand = np.arange(max(0, p_lo - 8 * nside), min(12 * nside * nside, p_hi + 8 * nside))
        dot = np.stack(ohp.pix2vec(nside, cand), axis=1) @ v
        inside, maybe = cand[dot > np.cos(r) + 4e-16], cand[dot > np.cos(r) - 4e-16]
        assert np.isin(inside, got).all() and np.isin(got, maybe).all(), (v, r, inside.size, got.size, maybe.size)
