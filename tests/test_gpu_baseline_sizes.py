"""GPU parity at the sizes BASELINE.json names ("Next round" item 1a).

The small-size parity tests (tests/test_gpu_parity.py) stop at NSIDE 256 / 256^2 / 64^3; these run the CUDA path
against the C oracle at the BASELINE configurations themselves:

* C1 in full (BaryonifyGrid / PaintProfilesGrid 2-D 512^2, 2 000 halos);
* C2 in full (BaryonifyShell NSIDE 1024, 10^5 halos, epsilon_max 20);
* the headline map, NSIDE 4096, on a sub-catalogue of the bench catalogue: the first 10^4 halos of the seed-42
  shuffle the reference's own driver uses (utils/Parallelize.py:255) + the 100 most massive + every halo whose
  disc reaches a pole -- BaryonifyShell and PaintProfilesShell;
* query_disc pixel lists bit-exact at NSIDE 1024 and 4096, and against a brute-force dot-product test that shares
  no code with either ring walk;
* BaryonifySnapshot at 128^3 particles / 10^4 halos with the PAIR LIST compared (sorted (particle, halo) ids,
  oracle side by brute force over all particles), not just the pair count.

Tolerances (north_star): indices / pair lists bit-exact, values <= 1e-5 relative (asserted far tighter), mass 1e-10.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MASS_RTOL = 1e-10


def relmax(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


def _threads():
    import os
    return os.cpu_count() or 1


# ------------------------------------------------------------------ C1
def test_c1_grid_2d_512_full(gpu):
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_grid()  # 512^2, 2 000 halos, L = 500/h Mpc, epsilon_max 20: BASELINE.json configs[0]
    assert cfg["field"].map.shape == (512, 512) and len(cfg["catalog"].cat) == 2000
    got_r = bf.BaryonifyGrid(cfg["catalog"], cfg["field"], 20.0, cfg["model"], verbose=False)
    ref_r = orc.BaryonifyGrid(cfg["catalog"], cfg["field"], 20.0, cfg["model"], nthreads=_threads())
    got, ref = got_r.process(), ref_r.process()
    assert got_r.npairs == ref_r.npairs and ref_r.npairs > 100_000
    assert relmax(got, ref) < 1e-9
    assert abs(got.sum() - cfg["field"].map.sum()) / cfg["field"].map.sum() < MASS_RTOL
    gp = bf.PaintProfilesGrid(cfg["catalog"], cfg["field"], 20.0, cfg["paint_model"], verbose=False)
    rp = orc.PaintProfilesGrid(cfg["catalog"], cfg["field"], 20.0, cfg["paint_model"], nthreads=_threads())
    g, r = gp.process(), rp.process()
    assert gp.npairs == rp.npairs
    assert np.array_equal(g != 0, r != 0)
    assert relmax(g, r) < 1e-11


# ------------------------------------------------------------------ C2
def test_c2_baryonify_shell_nside1024_full(gpu):
    import baryonification_b200 as bf
    from baryonification_b200 import _device as dv, synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_shell(nside=1024, nhalo=100_000)  # BASELINE.json configs[1]
    got_r = bf.BaryonifyShell(cfg["catalog"], cfg["field"], 20.0, cfg["model"], verbose=False)
    ref_r = orc.BaryonifyShell(cfg["catalog"], cfg["field"], 20.0, cfg["model"], nthreads=_threads())
    off_ref = ref_r.offsets()
    off_got = got_r.offsets_device(dv.device()).cpu().numpy()
    assert got_r.npairs == ref_r.npairs and ref_r.npairs > 5e7
    scale = np.max(np.abs(off_ref))
    assert np.max(np.abs(off_got - off_ref)) < 1e-13 * max(1.0, scale / 1e-3)
    del off_got
    ref = np.zeros(cfg["field"].map.size)
    orc._lib().bfo_regrid_shell(1024, np.ascontiguousarray(cfg["field"].map), off_ref, ref)
    got = got_r.process()
    assert relmax(got, ref) < 1e-9
    assert np.allclose(got, ref, rtol=1e-5, atol=1e-5 * ref.mean())
    assert abs(got.sum() - cfg["field"].map.sum()) / cfg["field"].map.sum() < MASS_RTOL


# ------------------------------------------------------------------ the headline map: NSIDE 4096
def _bench_subcatalogue(cfg, nside, first=10_000, massive=100):
    """First `first` halos of the seed-42 shuffle (Parallelize.py:255) + the `massive` most massive + every halo whose
    query disc reaches a pole, in catalogue order."""
    import copy
    from baryonification_b200 import host
    from baryonification_b200._ccl import default_mass_def
    cat = cfg["catalog"]
    n = len(cat.cat)
    s = host.shell_scalars(cat, cat.cosmology, default_mass_def(), cfg["epsilon_max"], cfg["model"], True)
    theta = np.arccos(np.clip(s["vec"][:, 2], -1, 1))
    polar = (theta - s["radius"] <= 0) | (theta + s["radius"] >= np.pi)
    idx = np.random.default_rng(42).choice(n, size=n, replace=False)[:first]
    keep = np.zeros(n, dtype=bool)
    keep[idx] = True
    keep[np.argsort(cat.cat["M"])[-massive:]] = True
    keep |= polar
    sub = copy.copy(cat)
    sub.cat = cat.cat[keep]
    return sub, int(polar.sum())


@pytest.fixture(scope="module")
def headline():
    from baryonification_b200 import synthetic as syn
    cfg = syn.config_shell(nside=4096, nhalo=1_000_000)  # bench.py's default workload (same seed)
    sub, npolar = _bench_subcatalogue(cfg, 4096)
    assert 10_100 <= len(sub.cat) <= 12_000 and npolar >= 1
    return cfg, sub


def test_headline_nside4096_subcatalogue_baryonify(gpu, headline):
    import baryonification_b200 as bf
    from baryonification_b200 import _device as dv
    from oracle import runners as orc
    cfg, sub = headline
    got_r = bf.BaryonifyShell(sub, cfg["field"], 20.0, cfg["model"], verbose=False)
    ref_r = orc.BaryonifyShell(sub, cfg["field"], 20.0, cfg["model"], nthreads=_threads())
    off_ref = ref_r.offsets()
    off_d = got_r.offsets_device(dv.device())
    assert got_r.npairs == ref_r.npairs and ref_r.npairs > 5e7
    scale = np.max(np.abs(off_ref))
    worst = 0.0
    step = 1 << 24  # compare in slices: the [Npix, 3] array is 4.8 GB
    for a in range(0, off_ref.shape[0], step):
        worst = max(worst, float(np.max(np.abs(off_d[a:a + step].cpu().numpy() - off_ref[a:a + step]))))
    assert worst < 1e-13 * max(1.0, scale / 1e-3), worst
    del off_d
    got = got_r.process()
    ref = np.zeros(cfg["field"].map.size)
    orc._lib().bfo_regrid_shell(4096, np.ascontiguousarray(cfg["field"].map), off_ref, ref)
    assert relmax(got, ref) < 1e-9
    assert abs(got.sum() - cfg["field"].map.sum()) / cfg["field"].map.sum() < MASS_RTOL


def test_headline_nside4096_subcatalogue_paint(gpu, headline):
    import copy
    import baryonification_b200 as bf
    from oracle import runners as orc
    cfg, sub = headline
    shell = copy.copy(cfg["field"])
    shell.map = np.zeros(cfg["field"].map.size)  # C3 paints onto a zero map
    got_r = bf.PaintProfilesShell(sub, shell, 20.0, cfg["paint_model"], verbose=False)
    ref_r = orc.PaintProfilesShell(sub, shell, 20.0, cfg["paint_model"], nthreads=_threads())
    got, ref = got_r.process(), ref_r.process()
    assert got_r.npairs == ref_r.npairs and ref_r.npairs > 5e7
    assert np.array_equal(got != 0, ref != 0)  # painted footprint = union of the pair lists
    assert relmax(got, ref) < 1e-11
    nz = ref != 0
    assert np.max(np.abs(got[nz] / ref[nz] - 1.0)) < 1e-9


# ------------------------------------------------------------------ query_disc at the large NSIDEs
@pytest.mark.parametrize("nside", [1024, 4096])
def test_healpix_query_disc_bit_exact_large_nside(gpu, nside):
    import torch
    from baryonification_b200 import _device as dv, _lib
    from oracle import healpix as ohp
    rng = np.random.default_rng(900 + nside)
    n = 120
    vec = rng.normal(size=(n, 3))
    vec /= np.linalg.norm(vec, axis=1)[:, None]
    vec[0], vec[1] = [0, 0, 1], [0, 0, -1]
    vec[2] = [1e-9, 0, 1]; vec[2] /= np.linalg.norm(vec[2])
    vec[3] = [1.0, -1e-12, 0.3]; vec[3] /= np.linalg.norm(vec[3])  # phi = 2 pi - eps: the ring wrap
    pixrad = np.sqrt(4 * np.pi / (12 * nside * nside))
    radius = pixrad * rng.choice([0.4, 1.0, 2.5, 7.0, 30.0, 150.0], size=n) * rng.uniform(0.8, 1.2, n)
    radius[:4] = [100 * pixrad, 40 * pixrad, 3 * pixrad, 60 * pixrad]
    for j in range(4, 14):  # discs that swallow a pole (centre at 0.7 r from it) or just touch it (1.0 r)
        d = radius[j] * (0.7 if j % 2 else 1.0)
        sgn = 1.0 if j % 4 < 2 else -1.0
        ph = rng.uniform(0, 2 * np.pi)
        vec[j] = [np.sin(d) * np.cos(ph), np.sin(d) * np.sin(ph), sgn * np.cos(d)]
    v, r = dv.to_dev(vec), dv.to_dev(radius)
    counts = dv.empty(n, dtype=torch.int64)
    _lib.call("bf_healpix_query_disc", nside, n, dv.ptr(v), dv.ptr(r), dv.ptr(counts), None, None, dv.stream_ptr())
    cnt = counts.cpu().numpy()
    offs = np.concatenate([[0], np.cumsum(cnt)])
    o = dv.to_dev(offs[:-1], dtype=np.int64)
    pix = dv.empty(int(offs[-1]) + 1, dtype=torch.int64)
    _lib.call("bf_healpix_query_disc", nside, n, dv.ptr(v), dv.ptr(r), dv.ptr(counts), dv.ptr(o), dv.ptr(pix),
              dv.stream_ptr())
    pix = pix.cpu().numpy()
    for j in range(n):
        ref = ohp.query_disc(nside, vec[j], radius[j])
        assert cnt[j] == ref.size, (j, cnt[j], ref.size)
        assert np.array_equal(pix[offs[j]:offs[j + 1]], ref), j
        # independent of both ring walks: every listed pixel centre is inside the disc (dot product), and the count
        # equals the brute-force count over the rings the disc can reach
        got = pix[offs[j]:offs[j + 1]]
        if got.size:
            pv = np.stack(ohp.pix2vec(nside, got), axis=1)
            ang = np.arccos(np.clip(pv @ vec[j], -1, 1))
            assert np.all(ang < radius[j] * (1 + 1e-9) + 1e-12), j
    # brute force on a few discs: all pixels of the bounding rings
    for j in (0, 1, 2, 3, 5, 6, 20, 50, 90):
        th0 = np.arccos(np.clip(vec[j, 2], -1, 1))
        t_lo, t_hi = max(th0 - radius[j] - 3 * pixrad, 0.0), min(th0 + radius[j] + 3 * pixrad, np.pi)
        p_lo = int(np.atleast_1d(ohp.ang2pix(nside, np.array([t_lo]), np.array([0.0])))[0])
        p_hi = int(np.atleast_1d(ohp.ang2pix(nside, np.array([t_hi]), np.array([0.0])))[0])
        p_lo, p_hi = max(0, p_lo - 8 * nside), min(12 * nside * nside, p_hi + 8 * nside)
        cand = np.arange(p_lo, p_hi)
        pv = np.stack(ohp.pix2vec(nside, cand), axis=1)
        dot = pv @ vec[j]
        inside = cand[dot > np.cos(radius[j]) + 4e-16]       # clearly inside
        maybe = cand[dot > np.cos(radius[j]) - 4e-16]        # inside or on the rounding edge
        got = set(pix[offs[j]:offs[j + 1]].tolist())
        assert set(inside.tolist()) <= got <= set(maybe.tolist()), j


def test_healpix_ang2vec_lonlat_kernel(gpu):
    """bf_healpix_ang2vec_lonlat (row a2 on the device) against the host routine the runners use and the oracle."""
    from baryonification_b200 import _device as dv, _lib, host
    from oracle import healpix as ohp
    rng = np.random.default_rng(5)
    n = 5000
    lon, lat = rng.uniform(0, 360, n), np.degrees(np.arcsin(rng.uniform(-1, 1, n)))
    lon[:4], lat[:4] = [0.0, 360.0, 180.0, 90.0], [90.0, -90.0, 0.0, 0.0]
    out = dv.empty((n, 3))
    lon_d, lat_d = dv.to_dev(lon), dv.to_dev(lat)
    _lib.call("bf_healpix_ang2vec_lonlat", n, dv.ptr(lon_d), dv.ptr(lat_d), dv.ptr(out), dv.stream_ptr())
    got = out.cpu().numpy()
    assert np.max(np.abs(got - host.ang2vec_lonlat(lon, lat))) < 4e-16
    assert np.max(np.abs(got - np.asarray(ohp.ang2vec(lon, lat, lonlat=True)).reshape(n, 3))) < 4e-16


# ------------------------------------------------------------------ snapshot: the pair LIST
def test_snapshot_pair_list_128cubed(gpu):
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_snapshot(npart_side=128, nhalo=10_000, dim=3, epsilon_max=5.0, L=300.0)
    got_r = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], 5.0, cfg["model"], verbose=False)
    ref_r = orc.BaryonifySnapshot(cfg["catalog"], cfg["field"], 5.0, cfg["model"], nthreads=_threads())
    got_pairs, ref_pairs = got_r.pair_list(), ref_r.pair_list()
    assert ref_pairs.size > 300_000
    assert np.array_equal(got_pairs, ref_pairs)  # (particle, halo) ids, bit-exact
    got, ref = got_r.process(), ref_r.process()
    assert got_r.npairs == ref_r.npairs == ref_pairs.size
    for k in "xyz":
        assert np.max(np.abs(got[k] - ref[k])) < 1e-11
    moved = np.zeros(got.size, dtype=bool)
    moved[(ref_pairs >> np.uint64(32)).astype(np.int64)] = True
    for k in "xyz":  # particles no halo reaches come back untouched
        assert np.array_equal(got[k][~moved], cfg["field"].cat[k][~moved])


@pytest.mark.parametrize("dim", [2, 3])
def test_snapshot_pair_list_small_and_wrapping(gpu, dim):
    """Query balls at the L/2 clip (they wrap the whole box) and 2-D snapshots: pair lists equal."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    side, eps, L = (300, 60.0, 300.0) if dim == 2 else (40, 60.0, 100.0)
    cfg = syn.config_snapshot(npart_side=side, nhalo=60, dim=dim, epsilon_max=eps, L=L)
    got_r = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], eps, cfg["model"], verbose=False)
    ref_r = orc.BaryonifySnapshot(cfg["catalog"], cfg["field"], eps, cfg["model"], nthreads=_threads())
    assert np.array_equal(got_r.pair_list(), ref_r.pair_list())


def test_snapshot_halo_grid_capacity_retry(gpu, monkeypatch):
    """>= 2^25 (cell, halo) entries on the proposed halo grid: the runner halves the grid until the list fits
    (BaryonifySnapshot.process() used to fail outright)."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_snapshot(npart_side=300, nhalo=60, dim=2, epsilon_max=60.0, L=300.0)
    monkeypatch.setenv("BF_SNAP_NG", "4096")  # 1.7e7 cells x 60 box-filling balls = 1e9 entries
    got_r = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], 60.0, cfg["model"], verbose=False)
    got = got_r.process()
    assert got_r.nentries < (1 << 25)
    ref_r = orc.BaryonifySnapshot(cfg["catalog"], cfg["field"], 60.0, cfg["model"], nthreads=_threads())
    ref = ref_r.process()
    assert got_r.npairs == ref_r.npairs
    for k in "xy":
        assert np.max(np.abs(got[k] - ref[k])) < 1e-9
