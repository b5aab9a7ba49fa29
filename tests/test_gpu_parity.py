"""GPU parity tests proper: CUDA path (through the C ABI) vs the CPU oracle on the same seeded inputs.

Tolerances (BASELINE.json north_star): pixel / cell indices and pair counts bit-exact; per-pixel values
within 1e-5 relative (we assert far tighter where the arithmetic allows); total mass within 1e-10.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

VAL_RTOL = 1e-5
MASS_RTOL = 1e-10


def relmax(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


# ------------------------------------------------------------------ HEALPix geometry
@pytest.mark.parametrize("nside", [1, 2, 8, 64, 1024, 4096])
def test_healpix_pix_roundtrip_and_vectors(gpu, nside):
    import torch
    from baryonification_b200 import _device as dv, _lib
    from oracle import healpix as ohp
    rng = np.random.default_rng(nside)
    npix = 12 * nside * nside
    pix = np.arange(npix) if npix <= 50000 else np.unique(np.concatenate(
        [rng.integers(0, npix, 40000), np.arange(2000), npix - 1 - np.arange(2000)]))
    p = dv.to_dev(pix, dtype=np.int64)
    th, ph, vec = dv.empty(pix.size), dv.empty(pix.size), dv.empty((pix.size, 3))
    _lib.call("bf_healpix_pix2ang", nside, pix.size, dv.ptr(p), dv.ptr(th), dv.ptr(ph), dv.stream_ptr())
    _lib.call("bf_healpix_pix2vec", nside, pix.size, dv.ptr(p), dv.ptr(vec), dv.stream_ptr())
    back = torch.empty_like(p)
    _lib.call("bf_healpix_ang2pix", nside, pix.size, dv.ptr(th), dv.ptr(ph), dv.ptr(back), dv.stream_ptr())
    assert np.array_equal(back.cpu().numpy(), pix)
    oth, oph = ohp.pix2ang(nside, pix)
    assert np.allclose(th.cpu().numpy(), oth, rtol=0, atol=4e-16 * np.pi)
    assert np.allclose(ph.cpu().numpy(), oph, rtol=0, atol=1e-15 * 2 * np.pi)
    ov = np.stack(ohp.pix2vec(nside, pix), axis=1)
    assert np.max(np.abs(vec.cpu().numpy() - ov)) < 1e-15


@pytest.mark.parametrize("nside", [4, 64, 1024])
def test_healpix_ang2pix_and_interp_weights(gpu, nside):
    from baryonification_b200 import _device as dv, _lib
    from oracle import healpix as ohp
    rng = np.random.default_rng(100 + nside)
    n = 20000
    lon = rng.uniform(0, 360, n)
    lat = np.degrees(np.arcsin(rng.uniform(-1, 1, n)))
    lat[:50] = np.linspace(-89.9999, 89.9999, 50)
    theta, phi = np.pi / 2 - np.radians(lat), np.radians(lon)
    t, p = dv.to_dev(theta), dv.to_dev(phi)
    pix = dv.empty(n, dtype=__import__("torch").int64)
    _lib.call("bf_healpix_ang2pix", nside, n, dv.ptr(t), dv.ptr(p), dv.ptr(pix), dv.stream_ptr())
    assert np.array_equal(pix.cpu().numpy(), ohp.ang2pix(nside, theta, phi))
    lo, la = dv.to_dev(lon), dv.to_dev(lat)
    ip, iw = dv.empty((n, 4), dtype=__import__("torch").int64), dv.empty((n, 4))
    _lib.call("bf_healpix_interp_weights_lonlat", nside, n, dv.ptr(lo), dv.ptr(la), dv.ptr(ip), dv.ptr(iw),
              dv.stream_ptr())
    op, ow = ohp.get_interp_weights(nside, lon, lat, lonlat=True)
    assert np.array_equal(ip.cpu().numpy(), op.T)
    assert np.max(np.abs(iw.cpu().numpy() - ow.T)) < 1e-9  # weights ~ 1e-16 * nside sensitivity
    assert np.allclose(iw.cpu().numpy().sum(1), 1.0, atol=1e-12)


@pytest.mark.parametrize("nside", [1, 4, 64, 512])
def test_healpix_query_disc_bit_exact(gpu, nside):
    import torch
    from baryonification_b200 import _device as dv, _lib
    from oracle import healpix as ohp
    rng = np.random.default_rng(200 + nside)
    n = 300
    vec = rng.normal(size=(n, 3))
    vec /= np.linalg.norm(vec, axis=1)[:, None]
    vec[0], vec[1] = [0, 0, 1], [0, 0, -1]
    vec[2] = [1e-9, 0, 1]; vec[2] /= np.linalg.norm(vec[2])
    radius = np.where(rng.random(n) < 0.8, rng.uniform(0, 0.2, n), rng.uniform(0, 3.3, n))
    radius[:3] = [0.3, 0.05, 0.1]
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


# ------------------------------------------------------------------ table readout
def test_table_readout_matches_oracle(gpu):
    from baryonification_b200 import _device as dv, synthetic as syn
    from baryonification_b200.tables import displacement_host_table, profile_host_table
    from oracle import runners as orc
    dm = syn.displacement_model(0.05, 1.2, 20.0)
    pm = syn.profile_model(0.05, 1.2, zero_beyond=50.0)
    rng = np.random.default_rng(7)
    r = np.concatenate([np.geomspace(5e-4, 5e2, 3000), np.exp(dm.raw_input_r_range[[0, 1, -2, -1]]), [0.0]])
    for M, z in [(3e13, 0.3), (1e15, 0.9), (2e12, 0.3), (1e14, 1.5), (float(np.exp(dm.raw_input_M_range[3])), 0.2)]:
        a = 1 / (1 + z)
        for model, mode, htab in ((dm, 0, displacement_host_table(dm)), (pm, 1, profile_host_table(pm, "2D"))):
            T = dv.DeviceTable(htab)
            rows = T.collapse(dv.to_dev([np.log(1 / a)]), dv.to_dev([np.log(M)]))
            R = model.mass_def.get_radius(model.cosmo, M, a) / a
            cut = 20.0 * R
            got = T.readout(rows[0], dv.to_dev(r), 0.0, cut, mode).cpu().numpy()
            want = orc.readout_displacement(dm, r, M, a) if mode == 0 else orc.readout_profile(pm, r, M, a)
            assert np.array_equal(np.isnan(got), np.isnan(want))
            ok = ~np.isnan(want)
            assert np.allclose(got[ok], want[ok], rtol=1e-12, atol=1e-300)


# ------------------------------------------------------------------ runners
@pytest.mark.parametrize("variant", ["gather", "atomic"])
def test_paint_shell_parity(gpu, variant):
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_shell(nside=128, nhalo=1500, zero_map=True, zero_beyond=50.0)
    got_r = bf.PaintProfilesShell(cfg["catalog"], cfg["field"], cfg["epsilon_max"], cfg["paint_model"], verbose=False,
                                  variant=variant)
    got = got_r.process()
    ref_r = orc.PaintProfilesShell(cfg["catalog"], cfg["field"], cfg["epsilon_max"], cfg["paint_model"])
    ref = ref_r.process()
    assert got_r.npairs == ref_r.npairs  # pair lists: same count ...
    assert np.array_equal(got != 0, ref != 0)  # ... and same footprint
    assert relmax(got, ref) < 1e-11
    assert np.allclose(got, ref, rtol=VAL_RTOL, atol=1e-30)


@pytest.mark.parametrize("variant", ["gather", "atomic"])
@pytest.mark.parametrize("nside,nhalo,eps", [(64, 600, 20.0), (256, 2500, 10.0), (8, 300, 30.0)])
def test_baryonify_shell_parity(gpu, nside, nhalo, eps, variant):
    import baryonification_b200 as bf
    from baryonification_b200 import _device as dv, synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_shell(nside=nside, nhalo=nhalo, epsilon_max=eps)
    got_r = bf.BaryonifyShell(cfg["catalog"], cfg["field"], eps, cfg["model"], verbose=False, variant=variant)
    ref_r = orc.BaryonifyShell(cfg["catalog"], cfg["field"], eps, cfg["model"])
    off_ref = ref_r.offsets()
    off_got = got_r.offsets_device(dv.device()).cpu().numpy()
    assert got_r.npairs == ref_r.npairs
    # (no exact-footprint check here: a zero displacement leaves normalise(pos) - vec = O(1e-17) rounding noise)
    assert np.max(np.abs(off_got - off_ref)) < 1e-13 * max(1.0, np.max(np.abs(off_ref)) / 1e-3)
    got, ref = got_r.process(), ref_r.process()
    assert np.allclose(got, ref, rtol=VAL_RTOL, atol=1e-5 * ref.mean())
    assert relmax(got, ref) < 1e-9
    assert abs(got.sum() - cfg["field"].map.sum()) / cfg["field"].map.sum() < MASS_RTOL


@pytest.mark.parametrize("variant", ["gather", "atomic"])
def test_baryonify_shell_tiny_halo_fallback(gpu, variant):
    """Discs holding < 4 pixel centres use the 4 interpolation neighbours (HealpixRunner.py:309-310)."""
    import baryonification_b200 as bf
    from baryonification_b200 import _device as dv, synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_shell(nside=16, nhalo=400, epsilon_max=1.0)
    got_r = bf.BaryonifyShell(cfg["catalog"], cfg["field"], 1.0, cfg["model"], verbose=False, variant=variant)
    ref_r = orc.BaryonifyShell(cfg["catalog"], cfg["field"], 1.0, cfg["model"])
    off_ref = ref_r.offsets()
    off_got = got_r.offsets_device(dv.device()).cpu().numpy()
    assert got_r.npairs == ref_r.npairs
    assert np.max(np.abs(off_got - off_ref)) < 1e-13


@pytest.mark.parametrize("dim,npix,nhalo,eps,ell", [(2, 256, 500, 20.0, False), (3, 64, 150, 8.0, False),
                                                     (2, 256, 500, 20.0, True)])
def test_grid_parity(gpu, dim, npix, nhalo, eps, ell):
    """ell: use_ellipticity = True with q_ell / A_ell columns (2-D only, Map2DRunner.py:525-530, :758-763)."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_grid(npix=npix, nhalo=nhalo, dim=dim, epsilon_max=eps, L=(500.0 / 0.7 if dim == 2 else 100.0),
                          ellipticity=ell)
    got_r = bf.BaryonifyGrid(cfg["catalog"], cfg["field"], eps, cfg["model"], ell, verbose=False)
    ref_r = orc.BaryonifyGrid(cfg["catalog"], cfg["field"], eps, cfg["model"], ell)
    got, ref = got_r.process(), ref_r.process()
    assert got_r.npairs == ref_r.npairs and ref_r.npairs > 1000
    assert got.shape == ref.shape
    assert relmax(got, ref) < 1e-9
    assert abs(got.sum() - cfg["field"].map.sum()) / cfg["field"].map.sum() < MASS_RTOL
    gp = bf.PaintProfilesGrid(cfg["catalog"], cfg["field"], eps, cfg["paint_model"], ell, verbose=False)
    rp = orc.PaintProfilesGrid(cfg["catalog"], cfg["field"], eps, cfg["paint_model"], ell)
    g, r = gp.process(), rp.process()
    assert gp.npairs == rp.npairs
    assert np.array_equal(g != 0, r != 0)
    assert relmax(g, r) < 1e-11


@pytest.mark.parametrize("layout", ["reference", "extra_column"])
@pytest.mark.parametrize("dim,side,nhalo,eps", [(3, 64, 400, 5.0), (2, 400, 300, 5.0), (3, 40, 60, 60.0)])
def test_snapshot_parity(gpu, dim, side, nhalo, eps, layout):
    """eps = 60 drives R_q to the L/2 clip (SnapshotRunner.py:215): query balls wrap the whole box."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_snapshot(npart_side=side, nhalo=nhalo, dim=dim, epsilon_max=eps, L=(100.0 if dim == 3 else 300.0))
    if layout == "extra_column":  # not the 32-byte (M, x, y, z) record: the field-by-field path
        old = cfg["field"].cat
        cat = np.zeros(old.size, dtype=old.dtype.descr + [("id", "i8")])
        for k in old.dtype.names:
            cat[k] = old[k]
        cat["id"] = np.arange(old.size)
        cfg["field"].cat = cat
    got_r = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], eps, cfg["model"], verbose=False)
    ref_r = orc.BaryonifySnapshot(cfg["catalog"], cfg["field"], eps, cfg["model"])
    got, ref = got_r.process(), ref_r.process()
    assert got_r.npairs == ref_r.npairs and ref_r.npairs > 100
    assert got.dtype == ref.dtype and got.shape == ref.shape
    for k in "xyz"[:dim]:
        assert np.max(np.abs(got[k] - ref[k])) < 1e-11
    for k in got.dtype.names:
        if k not in "xyz"[:dim]:
            assert np.array_equal(got[k], ref[k])
    again = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], eps, cfg["model"], verbose=False).process()
    assert all(np.array_equal(got[k], again[k]) for k in got.dtype.names)  # no atomics: bit-reproducible


def test_snapshot_dense_block_uses_several_queue_rounds(gpu):
    """More than SN_QCAP pairs inside one 256-particle block (particles piled onto a few halos)."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_snapshot(npart_side=24, nhalo=300, dim=3, epsilon_max=5.0, L=100.0)
    cat, hc = cfg["field"].cat, cfg["catalog"].cat
    rng = np.random.default_rng(11)
    for k in "xyz":  # all particles within 1.5 Mpc of the halo cloud's corner, halos crowded there too
        cat[k] = rng.uniform(10.0, 11.5, cat.size)
        hc[k] = rng.uniform(9.5, 12.0, hc.size).astype(hc.dtype[k])
    got_r = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], 5.0, cfg["model"], verbose=False)
    ref_r = orc.BaryonifySnapshot(cfg["catalog"], cfg["field"], 5.0, cfg["model"])
    got, ref = got_r.process(), ref_r.process()
    assert got_r.npairs == ref_r.npairs and ref_r.npairs > 50 * cat.size
    for k in "xyz":
        assert np.max(np.abs(got[k] - ref[k])) < 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("dim,n,ngrid", [(3, 300_000, 64), (2, 200_000, 256), (3, 5_000, 7)])
def test_make_map_equals_numpy_histogramdd_bit_for_bit(gpu, dim, n, ngrid):
    """ParticleSnapshot.make_map (utils/io.py:657-670) on the device: the bins of np.histogramdd (particles on cell
    edges, on the box edges x = 0 and x = L, outside the box) and np.bincount's summation order -> the same bits, for
    unequal masses too; bit-reproducible."""
    from baryonification_b200 import synthetic as syn
    rng = np.random.default_rng(100 + dim + ngrid)
    L = 250.0
    xyz = [rng.uniform(0.0, L, n) for _ in range(dim)]
    edges = np.linspace(0, L, ngrid + 1)
    for c in xyz:  # exact hits on interior edges, on both ends of the box, and strays outside it
        c[:200] = rng.choice(edges, 200)
        c[200:210] = L
        c[210:220] = 0.0
        c[220:230] = np.array([-1e-9, L + 1e-9, -5.0, 2 * L, np.nextafter(L, 0), np.nextafter(0, 1), np.nextafter(L, 2 * L),
                               -0.0, L / 3, 2 * L / 3])
        rng.shuffle(c)
    M = rng.lognormal(0.0, 2.0, n)  # unequal masses: the summation order matters
    snap = syn.ParticleSnapshot(x=xyz[0], y=xyz[1], z=xyz[2] if dim == 3 else None, M=M, L=L, redshift=0.0, cosmo=syn.COSMO)
    want = np.histogramdd(np.stack(xyz, axis=1), bins=(edges,) * dim, weights=M)[0]
    got = snap.make_map(ngrid)
    assert got.shape == want.shape and got.dtype == want.dtype
    assert np.array_equal(got, want)
    assert np.array_equal(got, snap.make_map(ngrid, device="host"))
    assert np.array_equal(got, snap.make_map(ngrid))
    wide = np.zeros(n, dtype=snap.cat.dtype.descr + [("id", "i8")])  # not the 32-byte record: the column path
    for k in snap.cat.dtype.names:
        wide[k] = snap.cat[k]
    snap.cat = wide
    assert np.array_equal(snap.make_map(ngrid), want)


@pytest.mark.gpu
def test_make_map_empty_and_nan_mass(gpu):
    from baryonification_b200 import synthetic as syn
    snap = syn.ParticleSnapshot(x=np.zeros(0), y=np.zeros(0), z=np.zeros(0), M=np.zeros(0), L=10.0, redshift=0.0, cosmo=syn.COSMO)
    assert np.array_equal(snap.make_map(8), np.zeros((8, 8, 8)))
    bad = syn.ParticleSnapshot(x=np.ones(3), y=np.ones(3), z=np.ones(3), M=np.nan, L=10.0, redshift=0.0, cosmo=syn.COSMO)
    with pytest.raises(AssertionError, match="provide a value for the particle mass"):
        bad.make_map(8)


@pytest.mark.gpu
def test_repeated_large_input_is_page_locked_in_place(gpu):
    """A >= 64 MiB map is staged through page-locked memory the first time it is processed and page-locked where it lies
    from the second time on (parameter sweeps process the same map again and again): same bits either way, and the
    registration ends with the array."""
    import gc
    import baryonification_b200 as bf
    from baryonification_b200 import _device as dv, synthetic as syn
    cfg = syn.config_shell(nside=1024, nhalo=3000)  # 12.6 Mpixel map = 96 MiB
    assert cfg["field"].map.nbytes >= dv._PIN_MIN_BYTES
    before = len(dv._pin_done)
    outs = []
    for _ in range(3):
        outs.append(bf.BaryonifyShell(cfg["catalog"], cfg["field"], cfg["epsilon_max"], cfg["model"], verbose=False).process())
    assert len(dv._pin_done) == before + 1
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])
    del cfg, outs
    gc.collect()
    assert len(dv._pin_done) == before


@pytest.mark.gpu
def test_snapshot_streaming_kernel_equals_chunked_kernel(gpu):
    """(M, x, y, z) records go through the TMA-fed streaming kernel, any other layout field by field through the chunked
    kernel: the same bits.  1.7e6 particles = more than one round of chunks per warp; the last tile is ragged."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    cfg = syn.config_snapshot(npart_side=120, nhalo=3000, dim=3, epsilon_max=5.0, L=200.0)
    rec = cfg["field"].cat[:-13].copy()
    cfg["field"].cat = rec
    a_r = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], 5.0, cfg["model"], verbose=False)
    a = a_r.process()
    wide = np.zeros(rec.size, dtype=rec.dtype.descr + [("id", "i8")])
    for k in rec.dtype.names:
        wide[k] = rec[k]
    cfg["field"].cat = wide
    b_r = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], 5.0, cfg["model"], verbose=False)
    b = b_r.process()
    assert a_r.npairs == b_r.npairs and a_r.npairs > 10_000
    for k in rec.dtype.names:
        assert np.array_equal(a[k], b[k]), k


@pytest.mark.gpu
def test_snapshot_chunked_transfer_pipeline_equals_one_shot(gpu, monkeypatch):
    """process() streams the records through HBM chunk by chunk (upload / kernel / read-back overlapped); ragged last
    chunk, more chunks than staging buffers, read-only input array."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from baryonification_b200.Runners import snapshot as snap
    cfg = syn.config_snapshot(npart_side=20, nhalo=120, dim=3, epsilon_max=5.0, L=100.0)
    want_r = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], 5.0, cfg["model"], verbose=False)
    want = want_r.process()
    monkeypatch.setattr(snap._RecordPipeline, "CHUNK", 999)  # 8000 particles -> 9 chunks, the last one short
    cfg["field"].cat.flags.writeable = False
    got_r = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], 5.0, cfg["model"], verbose=False)
    got = got_r.process()
    assert got_r.npairs == want_r.npairs
    for k in want.dtype.names:
        assert np.array_equal(got[k], want[k]), k


@pytest.mark.gpu
@pytest.mark.parametrize("cls", ["BaryonifyShell", "PaintProfilesShell"])
def test_shell_halo_chunk_pipeline_equals_one_pass(gpu, cls):
    """Large catalogues are evaluated in runs of consecutive halos (host set-up of run k+1 under the kernels of run k);
    same pair set, sums continue on the device in catalogue order (only the grouping of partial sums changes)."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    cfg = syn.config_shell(nside=128, nhalo=3001, zero_map=(cls == "PaintProfilesShell"))
    model = cfg["paint_model"] if cls == "PaintProfilesShell" else cfg["model"]
    one = getattr(bf, cls)(cfg["catalog"], cfg["field"], 20.0, model, verbose=False)
    want = one.process()
    piped = getattr(bf, cls)(cfg["catalog"], cfg["field"], 20.0, model, verbose=False)
    piped.pipeline_min_halos, piped.pipeline_chunks, piped.pipeline_upload_after = 1000, 3, 2
    got = piped.process()
    assert "pairs_pipelined" in piped.timings and "pairs_pipelined" not in one.timings
    assert piped.npairs == one.npairs
    assert relmax(got, want) < 1e-13
    if cls == "PaintProfilesShell":
        assert np.array_equal(got != 0, want != 0)
    again = getattr(bf, cls)(cfg["catalog"], cfg["field"], 20.0, model, verbose=False)
    again.pipeline_min_halos, again.pipeline_chunks = 1000, 3
    if cls == "PaintProfilesShell":
        assert np.array_equal(again.process(), got)  # still deterministic


# ------------------------------------------------------------------ empty / degenerate inputs
def _empty(cat):
    import copy
    c = copy.copy(cat)
    c.cat = cat.cat[:0]
    return c


@pytest.mark.gpu
def test_empty_halo_catalogue_for_every_runner(gpu):
    """The reference's loops simply do not run: Paint -> zeros, BaryonifyGrid / Snapshot -> the input, BaryonifyShell ->
    the input re-binned with zero offsets (interpolation weights of a pixel centre are 1 only to rounding)."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    sh = syn.config_shell(nside=16, nhalo=10)
    shz = syn.config_shell(nside=16, nhalo=10, zero_map=True)
    for variant in ("gather", "atomic"):
        got = bf.BaryonifyShell(_empty(sh["catalog"]), sh["field"], 20.0, sh["model"], verbose=False, variant=variant).process()
        want = orc.BaryonifyShell(_empty(sh["catalog"]), sh["field"], 20.0, sh["model"]).process()
        assert relmax(got, want) < 1e-13
        p = bf.PaintProfilesShell(_empty(shz["catalog"]), shz["field"], 20.0, shz["paint_model"], verbose=False, variant=variant)
        assert not p.process().any() and p.npairs == 0
    for dim in (2, 3):
        g = syn.config_grid(npix=16, nhalo=5, dim=dim, L=50.0)
        for variant in ("gather", "atomic"):
            b = bf.BaryonifyGrid(_empty(g["catalog"]), g["field"], 20.0, g["model"], verbose=False, variant=variant)
            assert np.array_equal(b.process(), g["field"].map) and b.npairs == 0
            assert not bf.PaintProfilesGrid(_empty(g["catalog"]), g["field"], 20.0, g["paint_model"], verbose=False,
                                            variant=variant).process().any()
        s = syn.config_snapshot(npart_side=6, nhalo=5, dim=dim, L=50.0)
        r = bf.BaryonifySnapshot(_empty(s["catalog"]), s["field"], 5.0, s["model"], verbose=False)
        out = r.process()
        assert r.npairs == 0 and all(np.array_equal(out[k], s["field"].cat[k]) for k in out.dtype.names)


@pytest.mark.gpu
def test_empty_particle_set_and_halos_outside_the_table(gpu):
    import copy
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    s = syn.config_snapshot(npart_side=8, nhalo=20, dim=3, L=50.0)
    none = copy.copy(s["field"])
    none.cat = s["field"].cat[:0]
    out = bf.BaryonifySnapshot(s["catalog"], none, 5.0, s["model"], verbose=False).process()
    assert out.shape == (0,) and out.dtype == s["field"].cat.dtype
    # masses far above the tabulated range: scipy's fill_value NaN -> the reference adds nothing for those halos
    g = syn.config_grid(npix=48, nhalo=40, dim=2, epsilon_max=20.0)
    g["catalog"].cat["M"][::2] *= 1e4
    for cls, model in ((bf.BaryonifyGrid, g["model"]), (bf.PaintProfilesGrid, g["paint_model"])):
        got = cls(g["catalog"], g["field"], 20.0, model, verbose=False).process()
        want = getattr(orc, cls.__name__)(g["catalog"], g["field"], 20.0, model).process()
        assert relmax(got, want) < 1e-11


# ------------------------------------------------------------------ committed reference fixtures
import cases  # noqa: E402


@pytest.mark.parametrize("name", cases.case_names())
def test_cuda_path_matches_reference_fixture(gpu, name):
    """CUDA path (through the C ABI) vs the outputs of the reference itself (tests/golden/*.npz)."""
    import baryonification_b200 as bf
    kind, cfg = cases.build(name)
    _, got = cases.run(bf, kind, cfg)
    cases.check(kind, got, cases.golden(name), cfg)


def test_cuda_readout_matches_reference_fixture(gpu):
    from baryonification_b200 import _device as dv, synthetic as syn
    from baryonification_b200.tables import displacement_host_table, profile_host_table
    g = cases.golden("readout")
    dm = syn.displacement_model(0.05, 1.2, 20.0)
    pm = syn.profile_model(0.05, 1.2, zero_beyond=50.0)
    r = dv.to_dev(g["r"])
    for i, (M, z) in enumerate(zip(g["M"], g["z"])):
        a = 1 / (1 + z)
        R = dm.mass_def.get_radius(dm.cosmo, M, a) / a
        for htab, mode, want in ((displacement_host_table(dm), 0, g["displacement"][i]),
                                 (profile_host_table(pm, "2D"), 1, g["projected"][i]),
                                 (profile_host_table(pm, "3D"), 1, g["real"][i])):
            T = dv.DeviceTable(htab)
            rows = T.collapse(dv.to_dev([np.log(1 / a)]), dv.to_dev([np.log(M)]))
            got = T.readout(rows[0], r, 0.0, 20.0 * R, mode).cpu().numpy()
            assert np.array_equal(np.isnan(got), np.isnan(want))
            ok = ~np.isnan(want)
            assert np.allclose(got[ok], want[ok], rtol=1e-12, atol=1e-300)


# ------------------------------------------------------------------ deterministic re-binning
def test_shell_regrid_sorted_is_bit_reproducible_and_matches_atomic(gpu):
    import torch
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    cfg = syn.config_shell(nside=128, nhalo=2000, epsilon_max=20.0)
    outs = {}
    for method in ("sorted", "sorted", "atomic"):
        r = bf.BaryonifyShell(cfg["catalog"], cfg["field"], 20.0, cfg["model"], verbose=False, regrid=method)
        outs.setdefault(method, []).append(r.process())
    a, b = outs["sorted"]
    assert np.array_equal(a, b)  # bit-identical run to run (gather + sorted re-binning: no atomics anywhere)
    assert np.allclose(a, outs["atomic"][0], rtol=1e-12, atol=0)
    tot = cfg["field"].map.sum()
    assert abs(a.sum() - tot) / tot < MASS_RTOL


def test_shell_regrid_sorted_equals_serial_order_sum(gpu):
    """With the SAME deposits (device weights), the sorted variant's per-pixel sums equal a serial (i, j)-ordered
    accumulation bit for bit -- the order of regrid_pixels_hpix (HealpixRunner.py:58-66)."""
    import torch
    from baryonification_b200 import _device as dv, _lib
    from oracle import healpix as ohp
    nside = 32
    npix = 12 * nside * nside
    rng = np.random.default_rng(3)
    orig = rng.lognormal(size=npix)
    orig[rng.random(npix) < 0.3] = 0.0
    off = rng.normal(scale=0.08, size=(npix, 3))
    od, fd, new = dv.to_dev(orig), dv.to_dev(off), dv.zeros(npix)
    lib = _lib.load()
    nb = lib.bf_shell_regrid_sorted_ws_bytes(nside, npix)
    ws = torch.empty(nb, dtype=torch.uint8, device=od.device)
    _lib.call("bf_shell_regrid_sorted", nside, 0, npix, dv.ptr(od), dv.ptr(fd), dv.ptr(new), dv.ptr(ws), nb,
              dv.stream_ptr())
    # device weights of the same displaced positions
    vec = np.stack(ohp.pix2vec(nside, np.arange(npix)), 1) + off
    dn = np.sqrt((vec ** 2).sum(1))
    theta, phi = np.arccos(vec[:, 2] / dn), np.arctan2(vec[:, 1], vec[:, 0])
    phi[phi < 0] += 2 * np.pi
    lo, la = dv.to_dev(np.degrees(phi)), dv.to_dev(90.0 - np.degrees(theta))
    ip, iw = dv.empty((npix, 4), dtype=torch.int64), dv.empty((npix, 4))
    _lib.call("bf_healpix_interp_weights_lonlat", nside, npix, dv.ptr(lo), dv.ptr(la), dv.ptr(ip), dv.ptr(iw),
              dv.stream_ptr())
    ip, iw = ip.cpu().numpy(), iw.cpu().numpy()
    want = np.zeros(npix)
    for i in np.flatnonzero(orig > 0):
        for j in range(4):
            want[ip[i, j]] += iw[i, j] * orig[i]
    got = new.cpu().numpy()
    # numpy's arccos / arctan2 and CUDA's differ by <= 2 ulp in the displaced angles; an interpolation weight
    # w = (phi - phi_i) / dphi close to 0 turns that into a large RELATIVE change of one tiny deposit, so the
    # comparison is relative to the map scale (the deposits' absolute error is ~1e-16 * nside * mass)
    assert np.max(np.abs(got - want)) < 1e-12 * np.max(want)
    # bit-exact wherever the device's own displaced angles agree with numpy's to the last bit (most pixels)
    assert np.mean(got == want) > 0.5


def _shell_regrid(kind, nside, orig, off):
    import ctypes as C
    import torch
    from baryonification_b200 import _device as dv, _lib
    lib = _lib.load(require_cuda=True)
    npix = 12 * nside * nside
    od, fd, new = dv.to_dev(orig), dv.to_dev(off), dv.zeros(npix)
    win = (C.c_int * 3)()
    if kind == "gather":
        nb = lib.bf_shell_regrid_gather_ws_bytes(nside)
        ws = torch.empty(nb, dtype=torch.uint8, device=od.device)
        rc = lib.bf_shell_regrid_gather(nside, dv.ptr(od), dv.ptr(fd), dv.ptr(new), dv.ptr(ws), nb, win, dv.stream_ptr())
    else:
        nb = lib.bf_shell_regrid_sorted_ws_bytes(nside, npix)
        ws = torch.empty(nb, dtype=torch.uint8, device=od.device)
        rc = lib.bf_shell_regrid_sorted(nside, 0, npix, dv.ptr(od), dv.ptr(fd), dv.ptr(new), dv.ptr(ws), nb, dv.stream_ptr())
    return rc, new.cpu().numpy(), tuple(win)


@pytest.mark.parametrize("nside,amp", [(1, 0.3), (2, 0.2), (4, 0.05), (16, 0.01), (16, 0.1), (64, 0.002), (64, 0.03),
                                       (256, 0.0005), (256, 0.004), (12, 0.02), (100, 0.003)])
def test_shell_regrid_gather_is_bit_identical_to_the_sorted_serial_order_sum(gpu, nside, amp):
    """Destination gather == counting sort bit for bit (same deposits, same ascending (source, j) order), from
    sub-pixel to many-pixel displacements, with masked pixels, on rings of every kind (caps, belt, poles)."""
    npix = 12 * nside * nside
    rng = np.random.default_rng(nside)
    orig = rng.lognormal(size=npix)
    orig[rng.random(npix) < 0.2] = 0.0
    off = rng.normal(scale=amp, size=(npix, 3))
    off[rng.random(npix) < 0.1] = 0.0
    rc_s, want, _ = _shell_regrid("sorted", nside, orig, off)
    rc_g, got, win = _shell_regrid("gather", nside, orig, off)
    assert rc_s == 0
    if rc_g == -3:  # BF_E_WINDOW: too wide for the gather at this amplitude; the runner then takes the sorted variant
        assert (win[1] - win[0] + 1) * (2 * win[2] + 1) > 165
        return
    assert rc_g == 0
    assert np.array_equal(got, want), (win, np.max(np.abs(got - want)))
    assert abs(got.sum() / orig.sum() - 1) < 1e-12


@pytest.mark.parametrize("nside,parts,amp", [(64, 3, 0.01), (32, 5, 0.03), (128, 2, 0.002)])
def test_shell_regrid_gather_partial_sums_of_ring_band_shards_add_up(gpu, nside, parts, amp):
    """The N-GPU form: every ring-band shard emits ITS sources only (keys around it are "no deposit") and gathers over
    the destinations they reach; the shards' partial maps add up to the whole-sphere result (bit-identical wherever a
    destination's sources lie in one shard, to rounding at the shard boundaries)."""
    import ctypes as C
    import torch
    from baryonification_b200 import _device as dv, _lib, sharding as sh
    lib = _lib.load(require_cuda=True)
    npix = 12 * nside * nside
    rng = np.random.default_rng(nside + parts)
    orig = rng.lognormal(size=npix)
    orig[rng.random(npix) < 0.2] = 0.0
    off = rng.normal(scale=amp, size=(npix, 3))
    rc, want, _ = _shell_regrid("gather", nside, orig, off)
    assert rc == 0
    nb = lib.bf_shell_regrid_gather_ws_bytes(nside)
    ws = torch.empty(nb, dtype=torch.uint8, device="cuda")
    ws.fill_(0x5a)  # stale keys everywhere: the margin must be rewritten by every shard
    total = np.zeros(npix)
    for lo, hi in sh.band_split(nside, parts):
        a, b = sh.band_pixel_range(nside, lo, hi)
        od, fd, new = dv.to_dev(orig[a:b]), dv.to_dev(off[a:b]), dv.zeros(npix)
        win = (C.c_int * 3)()
        _lib.check(lib.bf_shell_regrid_gather_emit_range(nside, a, b, dv.ptr(od), dv.ptr(fd), dv.ptr(ws), nb, win,
                                                         dv.stream_ptr()), "emit_range")
        reach = (max(win[1], -win[0]) + 1) * 4 * nside
        _lib.call("bf_shell_regrid_gather_range", nside, dv.ptr(ws), nb, win, max(0, a - reach), min(npix, b + reach),
                  dv.ptr(new), dv.stream_ptr())
        part = new.cpu().numpy()
        assert abs(part.sum() / orig[a:b].sum() - 1) < 1e-12  # a shard's partial map holds exactly its own mass
        total += part
    assert np.allclose(total, want, rtol=1e-13, atol=0)
    assert np.mean(total == want) > 0.9


@pytest.mark.parametrize("mode_name", ["paint", "baryonify"])
def test_band_shard_pair_stage_with_pipelined_read_back(gpu, mode_name):
    """The N-GPU paint path on one GPU: every ring-band shard evaluates its bands group by group and copies each finished
    group into ITS piece of one page-locked host array (``accumulate_to_host(out_h=...)``); the pieces assemble the
    whole-sphere accumulator bit for bit (pixel owners: no collective)."""
    import torch
    from baryonification_b200 import engine, host, sharding as sh, synthetic as syn
    from baryonification_b200._ccl import default_mass_def
    from baryonification_b200.tables import displacement_host_table, profile_host_table
    nside = 128
    cfg = syn.config_shell(nside=nside, nhalo=1500, zero_map=(mode_name == "paint"))
    bary = mode_name == "baryonify"
    model = cfg["model"] if bary else cfg["paint_model"]
    mode = engine.BARYONIFY if bary else engine.PAINT
    htab = displacement_host_table(model) if bary else profile_host_table(model, "2D")
    s = host.shell_scalars(cfg["catalog"], cfg["catalog"].cosmo, default_mass_def(), cfg["epsilon_max"], model, bary)
    whole = engine.ShellEngine(nside, mode, htab, s, gpu).upload().bin()
    want = whole.accumulate().cpu().numpy()
    npairs = whole.pair_count()
    shape = want.shape
    out_h = torch.empty(shape, dtype=torch.float64, pin_memory=True)
    out_h.fill_(float("nan"))
    total = 0
    for lo, hi in sh.band_split(nside, 3):
        eng = engine.ShellEngine(nside, mode, htab, s, gpu, bands=(lo, hi)).upload().bin()
        a, b = eng.pix_first, eng.pix_last
        eng.accumulate_to_host(chunks=4, out_h=out_h[a:b])
        total += eng.npairs
    assert total == npairs
    assert np.array_equal(out_h.numpy(), want)


def test_shell_regrid_gather_window_fallback_and_runner_default(gpu):
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    nside = 32
    npix = 12 * nside * nside
    rng = np.random.default_rng(5)
    rc, _, win = _shell_regrid("gather", nside, np.ones(npix), rng.normal(scale=0.5, size=(npix, 3)))
    assert rc == -3 and (win[1] - win[0] + 1) * (2 * win[2] + 1) > 165
    rc, out, win = _shell_regrid("gather", nside, np.zeros(npix), np.zeros((npix, 3)))  # nothing to deposit
    assert rc == 0 and not out.any()
    cfg = syn.config_shell(nside=64, nhalo=400)
    g = bf.BaryonifyShell(cfg["catalog"], cfg["field"], 20.0, cfg["model"], verbose=False)  # default: gather
    a = g.process()
    assert g.regrid_used == "gather"
    s = bf.BaryonifyShell(cfg["catalog"], cfg["field"], 20.0, cfg["model"], verbose=False, regrid="sorted")
    assert np.array_equal(a, s.process()) and s.regrid_used == "sorted"


def test_shell_regrid_sorted_bucket_overflow_path(gpu):
    """> 255 deposits into one destination (every pixel displaced onto the same point): the u8 ranks saturate, the
    device-side flag switches the bucket pass to its cursors (no host round trip), and the sums are still the serial
    entry-order sums -- bit-reproducible and equal to a host accumulation of the same deposits in (i, j) order."""
    import torch
    from baryonification_b200 import _device as dv, _lib
    from oracle import healpix as ohp
    nside = 16
    npix = 12 * nside * nside
    rng = np.random.default_rng(11)
    orig = rng.lognormal(size=npix)
    pv = np.stack(ohp.pix2vec(nside, np.arange(npix)), 1)
    target = np.array([0.3, -0.5, np.sqrt(1 - 0.34)])
    off = np.ascontiguousarray(target[None, :] - pv)  # every displaced vector is `target` (up to rounding)
    od, fd = dv.to_dev(orig), dv.to_dev(off)
    lib = _lib.load()
    nb = lib.bf_shell_regrid_sorted_ws_bytes(nside, npix)
    ws = torch.empty(nb, dtype=torch.uint8, device=od.device)
    outs = []
    for _ in range(2):
        new = dv.zeros(npix)
        _lib.call("bf_shell_regrid_sorted", nside, 0, npix, dv.ptr(od), dv.ptr(fd), dv.ptr(new), dv.ptr(ws), nb,
                  dv.stream_ptr())
        outs.append(new.cpu().numpy())
    assert np.array_equal(outs[0], outs[1])
    assert np.count_nonzero(outs[0]) <= 16 and abs(outs[0].sum() / orig.sum() - 1) < 1e-13
    at = dv.zeros(npix)
    _lib.call("bf_shell_regrid_range", nside, 0, npix, dv.ptr(od), dv.ptr(fd), dv.ptr(at), dv.stream_ptr())
    at = at.cpu().numpy()
    assert np.array_equal(at != 0, outs[0] != 0) and np.allclose(at, outs[0], rtol=1e-12, atol=0)


@pytest.mark.parametrize("dim,npix", [(2, 128), (3, 32)])
def test_grid_regrid_sorted_matches_atomic_and_oracle(gpu, dim, npix):
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_grid(npix=npix, nhalo=200, dim=dim, epsilon_max=10.0, L=(500.0 / 0.7 if dim == 2 else 100.0))
    s1 = bf.BaryonifyGrid(cfg["catalog"], cfg["field"], 10.0, cfg["model"], verbose=False, regrid="sorted").process()
    a1 = bf.BaryonifyGrid(cfg["catalog"], cfg["field"], 10.0, cfg["model"], verbose=False, regrid="atomic").process()
    ref = orc.BaryonifyGrid(cfg["catalog"], cfg["field"], 10.0, cfg["model"]).process()
    assert relmax(s1, a1) < 1e-12 and relmax(s1, ref) < 1e-9
    tot = cfg["field"].map.sum()
    assert abs(s1.sum() - tot) / tot < MASS_RTOL


@pytest.mark.gpu
@pytest.mark.parametrize("dim,npix,nhalo,eps,ell", [(2, 64, 120, 20.0, False), (2, 75, 200, 30.0, True), (2, 512, 2000, 20.0, False),
                                                    (3, 24, 40, 8.0, False), (3, 37, 120, 12.0, False), (3, 64, 300, 20.0, False)])
def test_grid_cell_owner_gather_equals_halo_centric_atomics(gpu, dim, npix, nhalo, eps, ell):
    """The two pair loops visit the same (halo, cell) set (exact pair count); values agree to rounding (different
    summation order, table-based ln); grids that are not a multiple of the tile; cut-outs wrapping the box."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    L = 500.0 / 0.7 if dim == 2 else 100.0
    cfg = syn.config_grid(npix=npix, nhalo=nhalo, dim=dim, epsilon_max=eps, L=L, ellipticity=ell)
    kw = dict(verbose=False, use_ellipticity=cfg.get("use_ellipticity", False))
    for cls, model in (("BaryonifyGrid", cfg["model"]), ("PaintProfilesGrid", cfg["paint_model"])):
        g = getattr(bf, cls)(cfg["catalog"], cfg["field"], eps, model, variant="gather", **kw)
        a = getattr(bf, cls)(cfg["catalog"], cfg["field"], eps, model, variant="atomic", **kw)
        got, want = g.process(), a.process()
        assert g.variant_used == "gather" and a.variant_used == "atomic"
        assert g.npairs == a.npairs and g.npairs > 500
        assert relmax(got, want) < 1e-11
        if cls == "PaintProfilesGrid":
            assert np.array_equal(got != 0, want != 0)
        again = getattr(bf, cls)(cfg["catalog"], cfg["field"], eps, model, variant="gather", **kw).process()
        assert np.array_equal(got, again)  # register sums in catalogue order: bit-reproducible


@pytest.mark.gpu
def test_grid_bin_refuses_a_cut_out_wider_than_the_grid(gpu):
    """Nsize > npix: numpy's fancy-index += drops the duplicate cells, which only the halo-centric path mimics; the
    runners never get there (Map2DRunner.py:489 / :733 clip the cut-out to half the box), the ABI says so."""
    import ctypes as C
    import torch
    from baryonification_b200 import _device as dv, _lib
    lib = _lib.load(require_cuda=True)
    t = {k: dv.to_dev(np.asarray(v, dtype=np.float64)) for k, v in dict(pos=[[3.0, 4.0]], cut=[1.0], shift=[0.0],
                                                                         rows=np.zeros((1, 8)), bins=np.arange(16.0)).items()}
    ns = dv.to_dev(np.array([40]), dtype=np.int64)
    desc = _lib.GridHalos(t["pos"].data_ptr(), ns.data_ptr(), t["cut"].data_ptr(), t["shift"].data_ptr(),
                          t["rows"].data_ptr(), None, 1)
    nb = lib.bf_grid_bin_ws_bytes(2, 16, 1, 64)
    ws = torch.empty(nb, dtype=torch.uint8, device=t["pos"].device)
    n = C.c_int64()
    rc = lib.bf_grid_bin_halos(2, 16, dv.ptr(t["bins"]), C.byref(desc), 8, 64, dv.ptr(ws), nb, C.byref(n), dv.stream_ptr())
    assert rc == -3 and b"wider than the grid" in lib.bf_last_error()


def _grid_regrid(kind, dim, npix, off, orig):
    """C-ABI call of one re-binning variant on host arrays -> (status, new map, window)."""
    import ctypes as C
    import torch
    from baryonification_b200 import _device as dv, _lib
    lib = _lib.load(require_cuda=True)
    off_d, orig_d = dv.to_dev(off), dv.to_dev(orig)
    new_d = dv.zeros(orig.size)
    win = (C.c_int * 3)()
    if kind == "gather":
        ws = torch.empty(lib.bf_grid_regrid_gather_ws_bytes(), dtype=torch.uint8, device=off_d.device)
        rc = lib.bf_grid_regrid_gather(dim, npix, dv.ptr(off_d), dv.ptr(orig_d), dv.ptr(new_d), dv.ptr(ws), ws.numel(),
                                       win, dv.stream_ptr())
    else:
        nb = lib.bf_grid_regrid_sorted_ws_bytes(dim, npix)
        ws = torch.empty(nb, dtype=torch.uint8, device=off_d.device)
        rc = lib.bf_grid_regrid_sorted(dim, npix, dv.ptr(off_d), dv.ptr(orig_d), dv.ptr(new_d), dv.ptr(ws), nb,
                                       dv.stream_ptr())
    torch.cuda.synchronize()
    return rc, new_d.cpu().numpy(), tuple(win)


@pytest.mark.gpu
@pytest.mark.parametrize("dim,npix,amp", [(2, 5, 0.9), (2, 37, 0.4), (2, 64, 2.7), (2, 100, 4.6), (3, 5, 0.6), (3, 19, 0.9),
                                          (3, 32, 1.7), (3, 21, 2.4), (2, 33, 0.0)])
def test_grid_regrid_gather_is_bit_identical_to_the_sorted_serial_order_sum(gpu, dim, npix, amp):
    """Destination gather == counting sort (both = the reference's serial summation order), for displacement windows of
    2..7 cells per axis, grids that are not a multiple of the tile, periodic wrap of the window, non-finite offsets and
    exact-integer displacements (zero-weight deposits)."""
    rng = np.random.default_rng(100 * dim + npix)
    ncell = npix ** dim
    off = rng.uniform(-amp, amp, (ncell, dim))
    off[rng.integers(0, ncell, 7)] = np.nan
    off[rng.integers(0, ncell, 7), 0] = np.inf
    off[rng.integers(0, ncell, 9)] = np.round(off[rng.integers(0, ncell, 9)])
    orig = rng.lognormal(0.0, 1.0, ncell)
    orig[rng.integers(0, ncell, 5)] = 0.0
    rc_s, want, _ = _grid_regrid("sorted", dim, npix, off, orig)
    rc_g, got, win = _grid_regrid("gather", dim, npix, off, orig)
    assert rc_s == 0 and rc_g == 0
    assert all(2 <= w <= 2 * int(np.ceil(amp)) + 2 for w in win[:dim])
    assert np.array_equal(got, want)
    assert abs(got.sum() - orig.sum()) / orig.sum() < 1e-13


@pytest.mark.gpu
def test_grid_regrid_gather_reports_a_too_wide_window_and_the_runner_falls_back(gpu):
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    rng = np.random.default_rng(5)
    off = rng.uniform(-9.0, 9.0, (40 * 40, 2))
    rc, _, win = _grid_regrid("gather", 2, 40, off, np.ones(1600))
    assert rc == -3 and win[0] * win[1] > 121  # BF_E_WINDOW
    cfg = syn.config_grid(npix=64, nhalo=120, dim=2, epsilon_max=20.0)
    g = bf.BaryonifyGrid(cfg["catalog"], cfg["field"], 20.0, cfg["model"], verbose=False)
    a = g.process()
    assert g.regrid_used == "gather" and max(g.regrid_window) <= 4
    s = bf.BaryonifyGrid(cfg["catalog"], cfg["field"], 20.0, cfg["model"], verbose=False, regrid="sorted").process()
    assert np.array_equal(a, s)


@pytest.mark.gpu
@pytest.mark.parametrize("dim,npix", [(2, 75), (3, 37)])
def test_grid_row_slabs_reassemble_the_single_call_result(gpu, dim, npix):
    """The slab entry points of the N > 1 path (bf_grid_gather_rows, bf_grid_regrid_window + _gather_rows), run slab
    after slab on one GPU with uneven, tile-unaligned slabs: same bits as the whole-grid calls."""
    import ctypes as C
    import torch
    import baryonification_b200 as bf
    from baryonification_b200 import _device as dv, _lib, sharding as sh, synthetic as syn
    lib = _lib.load(require_cuda=True)
    cfg = syn.config_grid(npix=npix, nhalo=150, dim=dim, epsilon_max=12.0, L=(500.0 / 0.7 if dim == 2 else 100.0))
    whole = bf.BaryonifyGrid(cfg["catalog"], cfg["field"], 12.0, cfg["model"], verbose=False)
    dev = dv.device(None)
    off_whole = whole.offsets_device(dev)
    slabs = [(0, 5), (5, 6), (6, npix - 9), (npix - 9, npix)]
    off = torch.zeros_like(off_whole)
    pairs = 0
    for a0, a1 in slabs:
        part = bf.BaryonifyGrid(cfg["catalog"], cfg["field"], 12.0, cfg["model"], verbose=False)
        part.rows = (a0, a1)
        o = part.offsets_device(dev)
        rl = npix ** (dim - 1) * dim
        off.view(npix, rl)[a0:a1] = o.view(npix, rl)[a0:a1]
        assert torch.count_nonzero(o.view(npix, rl)[:a0]) == 0 and torch.count_nonzero(o.view(npix, rl)[a1:]) == 0
        pairs += part.npairs
    assert pairs == whole.npairs and torch.equal(off, off_whole)
    off = off * 30.0  # exaggerate so that mass crosses slab boundaries
    orig = dv.to_dev(np.ascontiguousarray(cfg["field"].map).reshape(-1))
    ws = torch.empty(64, dtype=torch.uint8, device=dev)
    want = dv.zeros(orig.numel())
    win = (C.c_int * 3)()
    _lib.check(lib.bf_grid_regrid_gather(dim, npix, dv.ptr(off), dv.ptr(orig), dv.ptr(want), dv.ptr(ws), 64, win,
                                         dv.stream_ptr()), "whole")
    lo, hi = [10 ** 9] * 3, [-10 ** 9] * 3
    for a0, a1 in slabs:
        lohi = (C.c_int * 6)()
        _lib.call("bf_grid_regrid_window", dim, npix, dv.ptr(off), a0, a1, dv.ptr(ws), 64, lohi, dv.stream_ptr())
        lo = [min(a, b) for a, b in zip(lo, list(lohi)[:3])]
        hi = [max(a, b) for a, b in zip(hi, list(lohi)[3:])]
    lohi = (C.c_int * 6)(*(lo[:dim] + [0] * (3 - dim) + hi[:dim] + [0] * (3 - dim)))
    got = dv.zeros(orig.numel())
    for a0, a1 in slabs:
        w2 = (C.c_int * 3)()
        _lib.call("bf_grid_regrid_gather_rows", dim, npix, dv.ptr(off), dv.ptr(orig), dv.ptr(got), lohi, a0, a1, w2,
                  dv.stream_ptr())
        assert tuple(w2) == tuple(win)
    assert max(win) >= 3 and torch.equal(got, want)


# ------------------------------------------------------------------ property-keyed tables (p_keys)
def test_param_tabulated_profiles_parity(gpu):
    """ParamTabulatedProfile-style tables [Nz, NM, NR, Ncdelta] with a per-halo `cdelta` column
    (utils/Tabulate.py:524-621; HealpixRunner.py:407-441, Map2DRunner.py:707-801)."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    cfg = syn.config_shell(nside=128, nhalo=1500, zero_map=True, zero_beyond=50.0, p_keys=True)
    for variant in ("gather", "atomic"):
        g = bf.PaintProfilesShell(cfg["catalog"], cfg["field"], 20.0, cfg["paint_model"], verbose=False, variant=variant)
        r = orc.PaintProfilesShell(cfg["catalog"], cfg["field"], 20.0, cfg["paint_model"])
        got, ref = g.process(), r.process()
        assert g.npairs == r.npairs
        assert np.array_equal(got != 0, ref != 0)
        assert relmax(got, ref) < 1e-11
    for dim, npix, L in ((2, 128, 500.0 / 0.7), (3, 32, 100.0)):
        cg = syn.config_grid(npix=npix, nhalo=200, dim=dim, epsilon_max=8.0, L=L, p_keys=True)
        g = bf.PaintProfilesGrid(cg["catalog"], cg["field"], 8.0, cg["paint_model"], verbose=False)
        r = orc.PaintProfilesGrid(cg["catalog"], cg["field"], 8.0, cg["paint_model"])
        got, ref = g.process(), r.process()
        assert g.npairs == r.npairs and np.array_equal(got != 0, ref != 0)
        assert relmax(got, ref) < 1e-11
    # the reference's Baryonify runners refuse property keys unless the model is a ParamTabulatedProfile
    dm = cfg["model"]
    dm.p_keys = ["cdelta"]
    with pytest.raises(AssertionError, match="You asked to use"):
        bf.BaryonifyShell(cfg["catalog"], cfg["field"], 20.0, dm, verbose=False).process()
    dm.p_keys = []


def test_paint_grid_pipelined_readback_equals_single_launch(gpu):
    """Maps of >= 64 MB are painted in tile-aligned runs of rows whose read-back overlaps the next launch: the host
    map must be the single-launch device result bit for bit, with the same pair count."""
    import torch
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    cfg = syn.config_grid(npix=208, nhalo=400, dim=3, epsilon_max=8.0)
    assert cfg["field"].map.size >= (1 << 23)
    r = bf.PaintProfilesGrid(cfg["catalog"], cfg["field"], 8.0, cfg["paint_model"], verbose=False)
    got = r.process()
    n1 = r.npairs
    want = r.paint_device(torch.device("cuda", torch.cuda.current_device())).cpu().numpy().reshape(got.shape)
    assert r.npairs == n1 and np.array_equal(got, want) and np.count_nonzero(got) > 0


# ------------------------------------------------------------------ Delta P(k)
@pytest.mark.parametrize("ngrd,nk,factor", [(32, 12, 1), (48, 20, 2), (27, 9, 1), (64, 30, 8)])
def test_delta_pk_matches_notebook_recipe(gpu, ngrd, nk, factor):
    """NGP counts bit-exact, bin membership (mode counts, k centres) exact, P(k) to 1e-11 against the numpy
    restatement of examples/10_Reproduce_Schneider_deltaPk.ipynb."""
    from baryonification_b200.pk import PowerSpectrumGrid
    from oracle import pk as opk
    rng = np.random.default_rng(ngrd)
    L = 200.0
    pts = rng.uniform(0, L, (200_000, 3))
    pts[:1000] += rng.normal(scale=3.0, size=(1000, 3))  # some particles outside [0, L): the fold wraps them
    pts[0] = [0.0, L, L / factor]  # edges
    ps = PowerSpectrumGrid(ngrd, L, nk)
    grid = ps.density(pts, factor)
    k, P, nm = ps.power(grid, factor)
    ko, Po, kc, Map = opk.power(pts, ngrd, nk, L, factor)
    assert np.array_equal(grid.cpu().numpy(), Map.astype(np.float64))
    assert np.array_equal(nm, kc.astype(np.float64))
    ok = kc > 0
    assert np.allclose(k[ok], ko[ok], rtol=1e-13, atol=0)
    assert np.allclose(P[ok], Po[ok], rtol=1e-11, atol=0)
    k2, P2, _ = ps.power(grid, factor)
    assert np.array_equal(P, P2) and np.array_equal(k, k2)  # deterministic reduction: bit-reproducible


def test_delta_pk_matches_the_notebooks_own_code(gpu):
    """Against tests/golden/pk_notebook.npz -- outputs of the code cells of
    examples/10_Reproduce_Schneider_deltaPk.ipynb executed verbatim (tests/golden/make_golden_pk.py): NGP counts and mode
    counts bit-exact, the reported k centres (k_cen * factor, bins of the whole box) to 1e-13, P(k) to 1e-11."""
    import importlib.util
    import os
    from baryonification_b200.pk import PowerSpectrumGrid
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    spec = importlib.util.spec_from_file_location("make_golden_pk", os.path.join(here, "make_golden_pk.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    g = np.load(os.path.join(here, "pk_notebook.npz"))
    for name, seed, n, Ngrd, Nk, L, f in mg.CASES:
        pts = mg.particles(seed, n, L)
        ps = PowerSpectrumGrid(Ngrd, L, Nk)
        grid = ps.density(pts, f)
        k, P, nm = ps.power(grid, f)
        assert np.array_equal(grid.cpu().numpy(), g[name + ".Map"].astype(np.float64)), name
        kc = g[name + ".k_c"]
        assert np.array_equal(nm, kc.astype(np.float64)), name
        ok = kc > 0
        assert np.allclose(k[ok], (g[name + ".k_cen"] * f)[ok], rtol=1e-13, atol=0), name
        assert np.allclose(P[ok], g[name + ".Pk"][ok], rtol=1e-11, atol=0), name


@pytest.mark.parametrize("ngrd,nk,parts", [(32, 12, 2), (27, 9, 4), (20, 7, 8)])
def test_delta_pk_piecewise_binning_adds_up(gpu, ngrd, nk, parts):
    """bf_pk_bin_part over y-ranges of the spectrum (the pieces the slab-decomposed transform leaves on the ranks):
    mode counts add up exactly to the whole-spectrum binning, power and k sums to rounding; the 1-rank
    ShardedPowerSpectrum (2-D + 1-D transforms) agrees with PowerSpectrumGrid."""
    import ctypes as C
    import torch
    from baryonification_b200 import _device as dv, _lib, sharding as sh
    from baryonification_b200.pk import PowerSpectrumGrid
    rng = np.random.default_rng(ngrd)
    L = 150.0
    pts = rng.uniform(0, L, (100_000, 3))
    ps = PowerSpectrumGrid(ngrd, L, nk)
    grid = ps.density(pts)
    k, P, nm = ps.power(grid)
    _, val, k0, dk = ps._setup(1)
    F = torch.fft.rfftn(grid)
    lib = _lib.load(require_cuda=True)
    tot = np.zeros((3, nk))
    for j0, j1 in sh.index_split(ngrd, parts):
        piece = F[:, j0:j1].contiguous()
        ext = (C.c_int64 * 3)(ngrd, j1 - j0, ngrd // 2 + 1)
        off = (C.c_int64 * 3)(0, j0, 0)
        nb = lib.bf_pk_bin_part_ws_bytes(ext, nk)
        ws = torch.empty(max(nb, 256), dtype=torch.uint8, device=grid.device)
        out = dv.empty((3, nk), dev=grid.device)
        _lib.call("bf_pk_bin_part", ngrd, dv.ptr(piece) if piece.numel() else None, None, ext, off, val, k0, dk, nk,
                  dv.ptr(out), dv.ptr(ws), ws.numel(), dv.stream_ptr())
        tot += out.cpu().numpy()
    assert np.array_equal(tot[2], nm)
    ok = nm > 0
    assert np.allclose(tot[0][ok] / tot[2][ok], P[ok], rtol=1e-12, atol=0)
    assert np.allclose(tot[1][ok] / tot[2][ok], k[ok], rtol=1e-13, atol=0)
    sp = sh.ShardedPowerSpectrum(ngrd, L, nk)
    k1, P1, nm1 = sp.power(pts)
    assert np.array_equal(nm1, nm)
    assert np.allclose(P1[ok], P[ok], rtol=1e-11, atol=0) and np.allclose(k1[ok], k[ok], rtol=1e-13, atol=0)


def test_delta_pk_suppression_of_a_baryonified_snapshot(gpu):
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    from baryonification_b200.pk import PowerSpectrumGrid
    from oracle import pk as opk
    cfg = syn.config_snapshot(npart_side=64, nhalo=2000, dim=3, epsilon_max=5.0, L=200.0)
    new = bf.BaryonifySnapshot(cfg["catalog"], cfg["field"], 5.0, cfg["model"], verbose=False).process()
    ps = PowerSpectrumGrid(32, 200.0, 14)
    res = ps.suppression(new, cfg["field"].cat, factors=(1, 4))
    for f in (1, 4):
        B = np.stack([new[k] for k in "xyz"], 1)
        D = np.stack([cfg["field"].cat[k] for k in "xyz"], 1)
        k, PB, kc, _ = opk.power(B, 32, 14, 200.0, f)
        _, PD, _, _ = opk.power(D, 32, 14, 200.0, f)
        ok = kc > 0
        assert np.allclose(res[f"Factor{f}"]["S"][ok], (PB / PD)[ok], rtol=1e-10)
        assert np.allclose(res[f"Factor{f}"]["k"][ok], k[ok], rtol=1e-13)


@pytest.mark.gpu
def test_paint_split_join_sum_equals_one_pass(gpu):
    """The SplitJoinParallel property (utils/Parallelize.py:237-275, :316-318): shuffle the catalogue with seed 42, cut
    it into njobs runs of ceil(N / njobs) halos, paint each onto a zero map with a re-instantiated Runner (positional
    arguments, verbose=False) and np.sum the maps == one pass over the whole catalogue."""
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    cfg = syn.config_shell(nside=64, nhalo=1501, zero_map=True)
    runner = bf.PaintProfilesShell(cfg["catalog"], cfg["field"], 20.0, cfg["paint_model"], verbose=False)
    want = runner.process()
    njobs, n = 4, len(cfg["catalog"].cat)
    per = int(np.ceil(n / njobs))
    shuffled = runner.HaloLightConeCatalog[np.random.default_rng(42).choice(n, size=n, replace=False)]
    empty = type(runner.LightconeShell)(map=np.zeros_like(runner.LightconeShell.map), cosmo=runner.cosmo)
    outs, pairs = [], 0
    for i in range(njobs):
        sub = type(runner)(shuffled[i * per:(i + 1) * per], empty, runner.epsilon_max, runner.model,
                           runner.use_ellipticity, runner.mass_def, verbose=False)
        outs.append(sub.process())
        pairs += sub.npairs
    got = np.sum(outs, axis=0)
    assert pairs == runner.npairs
    assert np.array_equal(got != 0, want != 0)
    assert relmax(got, want) < 1e-13


@pytest.mark.gpu
@pytest.mark.parametrize("cls", ["BaryonifyShell", "PaintProfilesShell"])
def test_shell_setup_in_runs_equals_one_piece(gpu, cls, monkeypatch):
    """A repeated process() knows the size of the (cell, halo) key list from the previous call: the catalogue is then
    prepared on the host in runs of consecutive halos, and the upload, (z, M) collapse and binning of a finished run
    overlap the set-up of the next (bf_shell_bin_begin / _add / _end).  Same keys after the sort, hence the same bits."""
    import baryonification_b200 as bf
    from baryonification_b200 import host, synthetic as syn
    monkeypatch.setattr(host, "_PAR_MIN", 500)
    monkeypatch.setenv("BF_HOST_THREADS", "3")
    cfg = syn.config_shell(nside=128, nhalo=3001, zero_map=(cls == "PaintProfilesShell"))
    model = cfg["paint_model"] if cls == "PaintProfilesShell" else cfg["model"]
    r = getattr(bf, cls)(cfg["catalog"], cfg["field"], 20.0, model, verbose=False)
    r.SETUP_RUNS_MIN_HALOS, r.SETUP_RUNS = 1000, 4
    seen = []
    from baryonification_b200 import engine
    orig_add = engine.ShellEngine.add_run
    monkeypatch.setattr(engine.ShellEngine, "add_run", lambda self, j0, j1, pins: (seen.append((j0, j1)), orig_add(self, j0, j1, pins))[1])
    first = r.process()
    assert seen == []  # no size hint yet: one piece, with the counting pass
    second = r.process()
    assert [b - a for a, b in seen] and seen[0][0] == 0 and seen[-1][1] == 3001 and len(seen) == 4
    assert np.array_equal(first, second)
    r._bin_hint = {k: 10 for k in r._bin_hint}  # a hint that is too small: the binning is redone with the exact size
    assert np.array_equal(r.process(), first)
