"""CPU: the N > 1 host logic of baryonification_b200.sharding with world_size-2/3 gloo process groups.

The kernels cannot run here (no GPU, and the product has no CPU fallback), so every rank produces its LOCAL
stage results with the CPU oracle (test infrastructure) and the product's partitioning + collectives
(``band_split``, ``halos_touching_bands``, ``exchange_boundary``, ``gather_slices``, ``index_split``) turn them into
the global answer, which must equal the single-process oracle result."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from baryonification_b200 import sharding as sh, synthetic as syn


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(fn, world, *args):
    port = _free_port()
    mp.spawn(_entry, args=(world, port, fn, args), nprocs=world, join=True)


def _entry(rank, world, port, fn, args):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import ref_shim
        ref_shim.install()
        fn(rank, world, *args)
    finally:
        dist.destroy_process_group()


# ------------------------------------------------------------------ partitions (single process)
@pytest.mark.parametrize("nside,world", [(1, 1), (4, 2), (64, 3), (4096, 8), (1024, 5)])
def test_band_split_covers_sphere_with_balanced_pixel_counts(nside, world):
    bands = sh.band_split(nside, world)
    assert bands[0][0] == 0 and all(bands[i][1] == bands[i + 1][0] for i in range(world - 1))
    ranges = [sh.band_pixel_range(nside, *b) for b in bands]
    assert ranges[0][0] == 0 and ranges[-1][1] == 12 * nside * nside
    assert all(ranges[i][1] == ranges[i + 1][0] and ranges[i][1] > ranges[i][0] for i in range(world - 1))
    if nside >= 64:
        n = np.array([b - a for a, b in ranges], dtype=float)
        assert n.max() / n.min() < 1.2


def test_touched_extent_matches_brute_force_without_host_sync():
    """[lo, hi, widen] of the device-side extent search: equal to a numpy scan; ``widen`` exactly when the window's far
    edge is non-zero (the whole array must then be scanned)."""
    rng = np.random.default_rng(4)
    n = 400
    for trial in range(200):
        first, last = sorted(rng.integers(0, n + 1, 2))
        acc = np.zeros(n)
        k = rng.integers(0, 6)
        acc[rng.integers(0, n, k)] = rng.normal(size=k)
        acc[first:last] = 1.0
        window = [None, 1, 7, 50, 1000][trial % 5]
        lo, hi, widen = sh._touched_extent(torch.from_numpy(acc), int(first), int(last), window).tolist()
        a = 0 if window is None else max(0, first - window)
        b = n if window is None else min(n, last + window)
        below, above = np.flatnonzero(acc[a:first]) + a, np.flatnonzero(acc[last:b]) + last
        assert lo == (below[0] if below.size else first) and hi == (above[-1] + 1 if above.size else last)
        want_widen = (below.size > 0 and below[0] == a and a > 0) or (above.size > 0 and above[-1] == b - 1 and b < n)
        assert bool(widen) == bool(want_widen)


def test_index_split():
    assert sh.index_split(10, 3) == [(0, 3), (3, 6), (6, 10)]
    assert sh.index_split(2, 4) == [(0, 0), (0, 1), (1, 1), (1, 2)]


@pytest.mark.parametrize("nside,world,eps", [(32, 3, 20.0), (16, 2, 1.0), (64, 5, 40.0)])
def test_halo_replication_mask_never_drops_a_contributing_halo(nside, world, eps):
    from oracle import healpix as ohp, runners as orc
    cfg = syn.config_shell(nside=nside, nhalo=400, epsilon_max=eps)
    o = orc.BaryonifyShell(cfg["catalog"], cfg["field"], eps, cfg["model"])._halo_scalars()
    bands = sh.band_split(nside, world)
    kept_total = 0
    for lo, hi in bands:
        first, last = sh.band_pixel_range(nside, lo, hi)
        keep = sh.halos_touching_bands(o["vec"][:, 2], o["radius"], nside, lo, hi)
        kept_total += keep.sum()
        for j in np.flatnonzero(~keep):
            pix = ohp.query_disc(nside, o["vec"][j], o["radius"][j])
            if pix.size < 4:  # HealpixRunner.py:309-310
                pix = ohp.get_interp_weights(nside, cfg["catalog"].cat["ra"][j], cfg["catalog"].cat["dec"][j], lonlat=True)[0]
            assert not np.any((pix >= first) & (pix < last)), j
    if eps <= 20 and nside >= 32:
        assert kept_total < world * 400  # and it does prune


# ------------------------------------------------------------------ collectives under gloo
def _baryonify_ranks(rank, world, nside, eps):
    from oracle import runners as orc
    cfg = syn.config_shell(nside=nside, nhalo=300, epsilon_max=eps)
    ref_r = orc.BaryonifyShell(cfg["catalog"], cfg["field"], eps, cfg["model"])
    off = ref_r.offsets() * 40.0  # exaggerate the displacements so that mass crosses the shard boundaries
    orig = np.ascontiguousarray(cfg["field"].map)
    npix = orig.size
    want = np.zeros(npix)
    orc._lib().bfo_regrid_shell_range(nside, 0, npix, orig, off, want)
    bands = sh.band_split(nside, world)
    ranges = [sh.band_pixel_range(nside, *b) for b in bands]
    first, last = ranges[rank]
    local = np.zeros(npix)
    orc._lib().bfo_regrid_shell_range(nside, first, last, orig, off, local)  # this rank's own pixels only
    crossing = torch.tensor([np.count_nonzero(local[:first]) + np.count_nonzero(local[last:])])
    dist.all_reduce(crossing)
    assert int(crossing) > 0  # some mass does cross a shard boundary, so the exchange is exercised
    acc = torch.from_numpy(local)
    own = sh.exchange_boundary(acc, ranges, window=(4 * nside if rank == 0 else None))  # tiny window: falls back
    assert own.shape[0] == last - first
    full = sh.gather_slices(own, ranges).numpy()
    # a pixel gets (own deposits) + (ghost deposits): two partial sums instead of one serial sum -> last-bit changes only
    assert np.allclose(full, want, rtol=1e-13, atol=1e-13 * want.max())
    assert abs(full.sum() - orig.sum()) / orig.sum() < 1e-12
    sums = torch.tensor([float(own.sum())], dtype=torch.float64)
    dist.all_reduce(sums)
    assert abs(float(sums) - orig.sum()) / orig.sum() < 1e-12


@pytest.mark.parametrize("world,nside,eps", [(2, 16, 20.0), (3, 32, 20.0)])
def test_boundary_mass_exchange_and_gather_match_single_process(world, nside, eps):
    _run(_baryonify_ranks, world, nside, eps)


def _paint_ranks(rank, world, nside):
    from oracle import runners as orc
    cfg = syn.config_shell(nside=nside, nhalo=200, zero_map=True)
    want = orc.PaintProfilesShell(cfg["catalog"], cfg["field"], cfg["epsilon_max"], cfg["paint_model"]).process()
    ranges = [sh.band_pixel_range(nside, *b) for b in sh.band_split(nside, world)]
    first, last = ranges[rank]
    # pixel-owner sharding: the rank's slice is complete on its own (every halo that touches it is replicated to it)
    own = torch.from_numpy(want[first:last].copy())
    full = sh.gather_slices(own, ranges).numpy()
    assert np.array_equal(full, want)
    off3 = torch.from_numpy(np.stack([want[first:last]] * 3, axis=1).copy())  # [n, 3] slices gather too
    assert np.array_equal(sh.gather_slices(off3, ranges).numpy()[:, 1], want)
    # in place: the rank's slice already sits at its offset of the output (what the benchmark step does), ghosts around it
    buf = torch.full((want.size,), float("nan"), dtype=torch.float64)
    buf[first:last] = own
    out = sh.gather_slices(buf[first:last], ranges, out=buf)
    assert out.data_ptr() == buf.data_ptr() and np.array_equal(buf.numpy(), want)


def test_pixel_owner_gather_is_exact():
    _run(_paint_ranks, 2, 16)


def _snapshot_ranks(rank, world):
    from oracle import runners as orc
    cfg = syn.config_snapshot(npart_side=12, nhalo=80, dim=3, epsilon_max=5.0, L=100.0)
    want = orc.BaryonifySnapshot(cfg["catalog"], cfg["field"], 5.0, cfg["model"]).process()

    class OracleRunner(orc.BaryonifySnapshot):  # stands in for the GPU runner: same constructor contract
        def __init__(self, cat, snap, eps, model, mass_def=None, verbose=True, device=None):
            super().__init__(cat, snap, eps, model, mass_def, verbose)

    r = OracleRunner(cfg["catalog"], cfg["field"], 5.0, cfg["model"])
    got_r = sh.ShardedSnapshot(r)
    got = got_r.process()
    assert got.dtype == want.dtype and got.shape == want.shape
    for k in want.dtype.names:
        assert np.array_equal(got[k], want[k]), k  # particle-centric: sharding by index changes nothing
    ref_pairs = orc.BaryonifySnapshot(cfg["catalog"], cfg["field"], 5.0, cfg["model"])
    ref_pairs.process()
    assert got_r.npairs == ref_pairs.npairs


def test_snapshot_index_sharding_is_exact():
    _run(_snapshot_ranks, 2)


# ------------------------------------------------------------------ periodic grids: row slabs
def test_halo_row_mask_is_a_superset_of_the_exact_cut_out_rows():
    from baryonification_b200 import host
    from baryonification_b200._ccl import default_mass_def
    cfg = syn.config_grid(npix=64, nhalo=300, dim=3, epsilon_max=12.0, L=100.0)
    G = cfg["field"]
    s = host.grid_scalars(cfg["catalog"], G, default_mass_def(), 12.0, cfg["model"], True)
    bins = np.asarray(G.bins, dtype=np.float64)
    cen = np.array([np.argmin(np.abs(bins - x)) for x in s["pos"][:, 0]])  # Map2DRunner.py:505
    for a0, a1 in sh.index_split(64, 5):
        keep = sh.halos_touching_rows(s["pos"][:, 0], bins, s["Nsize"], a0, a1)
        for j in np.flatnonzero(~keep):
            w = s["Nsize"][j] // 2
            rows_j = np.arange(cen[j] - w, cen[j] + w) % 64  # pick_indices (:394-423)
            assert not np.any((rows_j >= a0) & (rows_j < a1)), j
        assert keep.sum() < 300  # and it prunes


def test_ghost_rows():
    rows = sh.index_split(10, 3)  # (0,3) (3,6) (6,10)
    assert sh.ghost_rows(rows, 0, 2, 1, 10) == [8, 9, 3]
    assert sh.ghost_rows(rows, 2, 1, 2, 10) == [5, 0, 1]
    assert sh.ghost_rows([(0, 10)], 0, 3, 3, 10) == []          # one rank owns everything
    assert sorted(sh.ghost_rows(rows, 1, 20, 20, 10)) == [0, 1, 2, 6, 7, 8, 9]  # window wider than the box
    assert sh._runs([0, 1, 2, 5, 7, 8]) == [(0, 3), (5, 6), (7, 9)]


def _ghost_exchange_ranks(rank, world, n, below, above):
    rng = np.random.default_rng(3)
    truth = rng.normal(size=(n, 7, 2))
    rows = sh.index_split(n, world)
    a0, a1 = rows[rank]
    buf = torch.full((n, 7, 2), float("nan"), dtype=torch.float64)
    buf[a0:a1] = torch.from_numpy(truth[a0:a1])
    sh.exchange_rows(buf, rows, below, above)
    have = sorted(list(range(a0, a1)) + sh.ghost_rows(rows, rank, below, above, n))
    got = buf.numpy()
    assert np.array_equal(got[have], truth[have])
    rest = np.setdiff1d(np.arange(n), have)
    assert np.all(np.isnan(got[rest]))


@pytest.mark.parametrize("world,n,below,above", [(2, 16, 2, 1), (3, 11, 1, 3), (3, 7, 4, 4)])
def test_ghost_row_exchange_fills_exactly_the_needed_rows(world, n, below, above):
    _run(_ghost_exchange_ranks, world, n, below, above)


def test_ghost_width_from_the_displacement_window_covers_every_deposit():
    """Rows outside [a0 - hi - 1, a1 - 1 - lo] never deposit into rows [a0, a1): the formula ShardedGrid uses."""
    rng = np.random.default_rng(8)
    n = 40
    off = rng.uniform(-2.3, 1.4, n * n)          # offsets along the first grid axis, one per cell
    a = np.repeat(np.arange(n), n)
    pos = (off + a) % n
    base = np.floor(pos).astype(int) % n
    d = (base - a + n // 2) % n - n // 2         # min-image floor(position) - cell, as the window kernel measures it
    lo, hi = d.min(), d.max()
    below, above = max(0, hi + 1), max(0, -lo)
    for a0, a1 in sh.index_split(n, 3):
        dest_hit = ((base >= a0) & (base < a1)) | (((base + 1) % n >= a0) & ((base + 1) % n < a1))
        allowed = np.zeros(n, dtype=bool)
        allowed[[g % n for g in range(a0 - below, a1 + above)]] = True
        assert np.all(allowed[a[dest_hit]])


# ------------------------------------------------------------------ Delta P(k): slab-decomposed transform
def _bin_piece(piece, j0, N, Nk, L):
    """Test-side k-binning of a [N, j1 - j0, N/2 + 1] piece of rfftn (what bf_pk_bin_part does on the device): the
    notebook's bin index (cell 12) with every mirrored mode (0 < l < N/2) counted twice."""
    kf = 2 * np.pi / L
    kbins = np.linspace(kf, kf * N / 2, Nk + 1)
    klin = np.fft.fftfreq(N, 1 / kf / N)
    nh = N // 2 + 1
    nj = piece.shape[1]
    k = np.sqrt((klin[:, None, None] ** 2 + klin[None, None, :nh] ** 2) + klin[None, j0:j0 + nj, None] ** 2)
    kinds = np.floor((k - kbins[0]) / (kbins[1] - kbins[0])).astype(int)
    l = np.arange(nh)[None, None, :]
    wt = np.where((l > 0) & (l < N / 2), 2.0, 1.0) * np.ones_like(k)
    m = (kinds >= 0) & (kinds < Nk)
    pw = (piece.real ** 2 + piece.imag ** 2) * wt
    return np.stack([np.bincount(kinds[m], weights=pw[m], minlength=Nk), np.bincount(kinds[m], weights=wt[m], minlength=Nk)])


def _slab_fft_ranks(rank, world, N, Nk):
    from oracle import pk as opk
    rng = np.random.default_rng(77)
    pts = rng.uniform(0, 50.0, size=(4000, 3))
    rows = sh.index_split(N, world)
    mine = pts[rank::world]                       # any partition of the particles works
    full = torch.from_numpy(opk.histogram3d(mine, N, 0.0, 50.0).astype(np.float64))
    slab = sh.reduce_to_slab(full, rows)
    Map = opk.histogram3d(pts, N, 0.0, 50.0)
    a, b = rows[rank]
    assert np.array_equal(slab.numpy(), Map[a:b])
    piece, (j0, j1) = sh.slab_rfftn(slab, rows)
    want = np.fft.rfftn(Map)
    assert piece.shape == (N, j1 - j0, N // 2 + 1)
    scale = np.abs(want).max()
    assert np.max(np.abs(piece.numpy() - want[:, j0:j1])) < 1e-12 * scale
    sums = torch.from_numpy(_bin_piece(piece.numpy(), j0, N, Nk, 50.0))
    dist.all_reduce(sums)
    _, Pk, k_c, _ = opk.power(pts, N, Nk, 50.0)
    assert np.array_equal(sums[1].numpy(), k_c)   # every mode of the cube exactly once across the pieces
    ok = k_c > 0
    assert np.allclose(sums[0].numpy()[ok] / k_c[ok], Pk[ok], rtol=1e-11)


@pytest.mark.parametrize("world,N,Nk", [(2, 16, 7), (3, 10, 5), (3, 9, 4), (2, 2, 1)])
def test_slab_decomposed_rfftn_and_piecewise_binning(world, N, Nk):
    _run(_slab_fft_ranks, world, N, Nk)


# ------------------------------------------------------------------ round 2: fixed-margin exchange, shared host arena
def _margin_ranks(rank, world, n, F):
    """Every rank deposits random mass over its range widened by F; the fixed-margin exchange must give every owner the
    sum of all deposits on its pixels (what exchange_boundary gives with its negotiated extents)."""
    rng = np.random.default_rng(5)
    ranges = sh.index_split(n, world)
    assert sh.fixed_margins_ok(ranges, F)
    want = np.zeros(n)
    locals_ = []
    for r, (a, b) in enumerate(ranges):
        w0, w1 = a - (F if r > 0 else 0), b + (F if r < world - 1 else 0)
        dep = rng.random(w1 - w0)
        dep[rng.random(w1 - w0) < 0.5] = 0.0
        want[w0:w1] += dep
        locals_.append(dep)
    own = sh.exchange_margins(torch.from_numpy(locals_[rank].copy()), ranges, F)
    a, b = ranges[rank]
    assert own.numel() == b - a
    assert np.allclose(own.numpy(), want[a:b], rtol=1e-15, atol=0)
    # ... and the shared host arena: every rank writes its piece, every rank sees the whole array
    arena = sh.SharedHostArena(n)
    arena.tensor[a:b].copy_(own)
    dist.barrier()
    assert np.allclose(arena.array, want, rtol=1e-15, atol=0)
    again = sh.SharedHostArena(n)  # a second arena is a different segment
    again.array[a:b] = -1.0
    dist.barrier()
    assert np.allclose(arena.array, want, rtol=1e-15, atol=0) and np.all(again.array == -1.0)


@pytest.mark.parametrize("world,n,F", [(2, 1000, 100), (3, 999, 333), (3, 40, 7)])
def test_fixed_margin_exchange_and_shared_host_arena(world, n, F):
    _run(_margin_ranks, world, n, F)


def test_fixed_margins_need_bands_longer_than_the_margin():
    assert not sh.fixed_margins_ok([(0, 100)], 10)                   # one rank: nothing to exchange
    assert not sh.fixed_margins_ok([(0, 100), (100, 105)], 10)       # a range shorter than the margin
    assert sh.fixed_margins_ok([(0, 100), (100, 110)], 10)
    ranges = [sh.band_pixel_range(4096, *b) for b in sh.band_split(4096, 8)]
    assert sh.fixed_margins_ok(ranges, sh.margin_pixels(4096))       # the benchmark partition takes the fast path
    small = [sh.band_pixel_range(16, *b) for b in sh.band_split(16, 2)]
    assert not sh.fixed_margins_ok(small, sh.margin_pixels(16))      # small spheres: the general exchange


# ------------------------------------------------------------------ split host set-up of the shell runners
def _w_split_setup(rank, world, nside, nhalo, baryonify):
    import baryonification_b200 as bf
    from baryonification_b200 import host
    from baryonification_b200._ccl import default_mass_def
    cfg = syn.config_shell(nside=nside, nhalo=nhalo, zero_map=not baryonify)
    model = cfg["model"] if baryonify else cfg["paint_model"]
    cls = bf.BaryonifyShell if baryonify else bf.PaintProfilesShell
    runner = cls(cfg["catalog"], cfg["field"], cfg["epsilon_max"], model, verbose=False)
    shard = sh.ShardedShell(runner)
    bands, _ = shard.plan()
    got = shard.scalars_device(torch.device("cpu"), bands[rank])  # 1/N of the host work + one all-gather + band mask
    s = host.shell_scalars(cfg["catalog"], cfg["catalog"].cosmo, default_mass_def(), cfg["epsilon_max"], model, baryonify)
    keep = sh.halos_touching_bands(s["vec"][:, 2], s["radius"], nside, *bands[rank])
    # the device-side mask may keep a few more halos than the host form, never fewer, and always in catalogue order
    full_idx = np.flatnonzero(np.isin(s["lon"], got["lon"].numpy()) & np.isin(s["lat"], got["lat"].numpy()))
    assert got["a"].numel() >= int(keep.sum())
    for k in ("a", "D", "radius", "z_in", "M_in", "lon", "lat", "cut", "lnr_shift"):
        g = got[k].numpy()
        assert g.ndim == 1 and got[k].is_contiguous()
        sel = s[k][keep]
        pos = np.searchsorted(full_idx, np.flatnonzero(keep))  # the kept halos inside the (super)set this rank holds
        assert np.array_equal(g[pos], sel), k  # bit-identical to the one-process set-up
    assert got["vec"].shape == (got["a"].numel(), 3) and got["vec"].is_contiguous()
    pos = np.searchsorted(full_idx, np.flatnonzero(keep))
    assert np.array_equal(got["vec"].numpy()[pos], s["vec"][keep])
    assert np.all(np.diff(full_idx) > 0)
    assert got["extras"] is None


@pytest.mark.parametrize("world,nside,nhalo,baryonify", [(2, 32, 301, True), (3, 16, 200, False), (3, 8, 2, True)])
def test_split_host_setup_all_gathers_the_catalogue_in_order(world, nside, nhalo, baryonify):
    """ShardedShell.scalars_device: every rank prepares 1/N of the halos (uneven pieces, pieces of zero halos), one
    all-gather, band mask -- each rank ends up with the one-process per-halo scalars of (at least) the halos that can
    touch its bands, bit for bit and in catalogue order."""
    _run(_w_split_setup, world, nside, nhalo, baryonify)
