"""CPU: host-side logic of the product -- vectorised per-halo scalar set-up (bit-identical to the reference's
halo-by-halo expressions, restated in oracle/runners.py), table export, containers, error behaviour."""
import numpy as np
import pytest

from baryonification_b200 import host, synthetic as syn, tables
from baryonification_b200._ccl import default_mass_def


def test_shell_scalars_bit_identical_to_per_halo_loop():
    from oracle import runners as orc
    cfg = syn.config_shell(nside=16, nhalo=500)
    s = host.shell_scalars(cfg["catalog"], cfg["catalog"].cosmo, default_mass_def(), 20.0, cfg["model"], True)
    o = orc.BaryonifyShell(cfg["catalog"], cfg["field"], 20.0, cfg["model"])._halo_scalars()
    for k in ("a", "D", "radius", "z_in", "M_in", "lon", "lat"):
        assert np.array_equal(s[k], o[k]), k
    assert np.max(np.abs(s["vec"] - o["vec"])) < 3e-16  # numpy vs libm sin/cos, <= 1 ulp
    cut, shift = orc._model_cut(cfg["model"], o["M"], o["a"])
    assert np.array_equal(s["cut"], cut) and np.array_equal(s["lnr_shift"], shift)


@pytest.mark.parametrize("dim", [2, 3])
def test_grid_and_snapshot_scalars_match_per_halo_loop(dim):
    cfg = syn.config_grid(npix=64, nhalo=300, dim=dim, L=(500 / 0.7 if dim == 2 else 100.0))
    cat, G = cfg["catalog"], cfg["field"]
    md = default_mass_def()
    cosmo = host.runner_cosmology(cat.cosmology, with_w0=False)
    a = 1 / (1 + cat.redshift)
    for baryonify in (True, False):
        s = host.grid_scalars(cat, G, md, 20.0, cfg["model"], baryonify)
        for j in range(0, len(cat), 7):  # Map2DRunner.py:484-497 / :722-733
            R = md.get_radius(cosmo, cat.cat["M"][j], a)
            if baryonify:
                Ns = int((2 * np.clip(20.0 * R / a, 0, np.max(G.bins) / 2) / G.res) // 2) * 2
            else:
                Ns = np.clip(int((2 * 20.0 * (R / a) / G.res) // 2) * 2, 2, G.bins.size // 2)
                assert s["cut"][j] == (R / a) * 20.0
            assert s["Nsize"][j] == Ns
            assert s["M_in"][j] == np.float64(np.log(cat.cat["M"][j]))  # float32 log, widened
        assert s["pos"].shape == (len(cat), dim) and s["pos"].dtype == np.float64
    snap = syn.config_snapshot(npart_side=8, nhalo=100, dim=dim, L=100.0)
    ss = host.snapshot_scalars(snap["catalog"], snap["field"], md, 5.0, snap["model"])
    assert ss["Rq"].max() <= 50.0 and ss["dim"] == dim


def test_nd_catalog_is_big_endian_float32_like_the_reference():
    cfg = syn.config_grid(npix=16, nhalo=10)
    c = cfg["catalog"].cat
    assert c.dtype["M"].str == ">f4" and c.dtype["x"].str == ">f4"  # utils/io.py:204-205
    lc = syn.config_shell(nside=4, nhalo=10)["catalog"].cat
    assert lc.dtype["M"] == np.float64


def test_lightcone_catalog_pole_clip_and_indexing():
    import baryonification_b200 as bf
    cat = bf.HaloLightConeCatalog(ra=[0.0, 10.0, 20.0], dec=[90.0, -90.0, 5.0], M=[1e14] * 3, z=[0.1, 0.2, 0.3],
                                  cosmo=syn.COSMO)
    assert abs(cat.cat["dec"][0]) < 90 and abs(cat.cat["dec"][1]) < 90  # utils/io.py:65-68
    sub = cat[np.array([2, 0])]
    assert isinstance(sub, bf.HaloLightConeCatalog) and len(sub) == 2 and sub.cat["z"][0] == 0.3
    with pytest.raises(Exception):
        bf.HaloLightConeCatalog(ra=[0.0], dec=[0.0], M=[1e14], z=[0.1], cosmo={"Omega_m": 0.3})


def test_containers_match_reference_containers():
    from oracle import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("reference tree not present")
    import refmodels as rm
    bfg = ref_shim.import_reference()
    cfg = syn.config_grid(npix=16, nhalo=20, dim=3, L=50.0)
    ours, ref = cfg["catalog"], rm.ref_nd_catalog(bfg, cfg["catalog"], 3)
    assert ours.cat.dtype == ref.cat.dtype and ours.cat.tobytes() == ref.cat.tobytes()
    g, gr = cfg["field"], rm.ref_grid(bfg, cfg["field"])
    assert (g.Npix, g.res, g.is2D) == (gr.Npix, gr.res, gr.is2D)
    assert np.array_equal(g.inds, gr.inds)
    sc = syn.config_shell(nside=4, nhalo=20)
    assert sc["catalog"].cat.tobytes() == rm.ref_lightcone_catalog(bfg, sc["catalog"]).cat.tobytes()
    assert sc["field"].NSIDE == rm.ref_shell(bfg, sc["field"]).NSIDE
    sn = syn.config_snapshot(npart_side=4, nhalo=5, L=20.0)
    assert sn["field"].cat.tobytes() == rm.ref_snapshot(bfg, sn["field"]).cat.tobytes()


def test_table_export_and_errors():
    dm = syn.displacement_model(0.05, 1.2, 20.0)
    t = tables.displacement_host_table(dm)
    assert t.values.shape == (10, 10, 500) and not t.is_log
    pm = syn.profile_model(0.05, 1.2, zero_beyond=50.0)
    tp = tables.profile_host_table(pm, "2D")
    assert tp.is_log and np.isneginf(tp.values).any()  # ln 0 kept as -inf (Tabulate.py:237)

    class Analytic:  # an un-tabulated profile: no raw_input_* -> TypeError, never a CPU evaluation
        def projected(self, *a):
            return 0

    with pytest.raises(TypeError, match="no tabulated data"):
        tables.profile_host_table(Analytic(), "2D")

    class Untabulated:  # reference message when setup_interpolator() was not run (BaryonCorrection.py:426)
        def displacement(self, *a):
            return 0

    with pytest.raises(NameError, match="No Table created"):
        tables.displacement_host_table(Untabulated())
    with pytest.raises(ValueError):
        tables.HostTable([0, 1], [0, 1], [1, 0], np.zeros((2, 2, 2)), False)


def test_runner_constructor_contract():
    """Positional order and public attributes SplitJoinParallel relies on (utils/Parallelize.py:237-243,271)."""
    import baryonification_b200 as bf
    cfg = syn.config_shell(nside=4, nhalo=10)
    md = default_mass_def()
    r = bf.PaintProfilesShell(cfg["catalog"], cfg["field"], 20.0, cfg["paint_model"], False, md, False)
    for attr in ("HaloLightConeCatalog", "LightconeShell", "cosmo", "model", "mass_def", "epsilon_max",
                 "use_ellipticity", "verbose"):
        assert hasattr(r, attr)
    r2 = type(r)(r.HaloLightConeCatalog[np.arange(5)], r.LightconeShell, r.epsilon_max, r.model, r.use_ellipticity,
                 r.mass_def, verbose=False)
    assert len(r2.HaloLightConeCatalog) == 5
    with pytest.raises(NotImplementedError):  # HealpixRunner.py:164-165
        bf.BaryonifyShell(cfg["catalog"], cfg["field"], 20.0, cfg["model"], True)
    import pickle
    pickle.loads(pickle.dumps(r))  # runners stay picklable (joblib/loky), no live CUDA handles on self


def test_spline_eval_is_bit_identical_to_scipy_and_chunked_setup_to_serial(monkeypatch):
    from scipy import interpolate
    z_t = np.linspace(0, 30, 1000)
    cs = interpolate.CubicSpline(z_t, np.sqrt(z_t + 0.3) * 1000 / (1 + z_t))
    rng = np.random.default_rng(5)
    x = np.concatenate([rng.uniform(0, 30, 200_000), z_t, [0.0, 30.0, 31.0, -0.5]])
    assert np.array_equal(host.spline_eval(cs, x), cs(x))
    cfg = syn.config_shell(nside=8, nhalo=3000)
    args = (cfg["catalog"], cfg["catalog"].cosmo, default_mass_def(), 20.0, cfg["model"], True)
    serial = host.shell_scalars(*args)
    monkeypatch.setattr(host, "_PAR_MIN", 100)
    monkeypatch.setenv("BF_HOST_THREADS", "4")
    par = host.shell_scalars(*args)
    for k, v in serial.items():
        assert np.array_equal(v, par[k]) and par[k].flags.c_contiguous, k


def test_shear_matrices_equal_per_halo_build_Rmat_and_3d_raises():
    """host.shear_matrices (vectorised) == the reference's per-halo build_Rmat (Map2DRunner.py:275-344, restated in
    oracle.runners) bit for bit; 3-D ellipticity raises the reference's exceptions (:565 / :790)."""
    from oracle import runners as orc
    cfg = syn.config_grid(npix=32, nhalo=1500, dim=2, ellipticity=True)
    got = host.shear_matrices(cfg["catalog"].cat)
    want = orc.BaryonifyGrid(cfg["catalog"], cfg["field"], 20.0, cfg["model"], use_ellipticity=True)._rmat(True)
    assert np.array_equal(got, want)
    assert np.array_equal(got[0], [1.0, 0.0, 0.0, 1.0])  # q = 1: identity
    assert np.allclose(got[:, 0] * got[:, 3] - got[:, 1] * got[:, 2], 1.0, atol=1e-12)  # shear: unit determinant
    cfg3 = syn.config_grid(npix=16, nhalo=10, dim=3, L=100.0, ellipticity=True)
    with pytest.raises(NotImplementedError):
        host.grid_rmat(cfg3["catalog"], cfg3["field"], True, True)
    with pytest.raises(ValueError):
        host.grid_rmat(cfg3["catalog"], cfg3["field"], True, False)


def test_runners_are_picklable_like_the_reference_ones():
    """utils/Parallelize.py:105-107 ships Runner objects to loky workers: no live CUDA handle may sit on ``self``."""
    import pickle
    import baryonification_b200 as bf
    from baryonification_b200 import synthetic as syn
    sh, gr, sn = syn.config_shell(nside=8, nhalo=10), syn.config_grid(npix=16, nhalo=5, dim=2), \
        syn.config_snapshot(npart_side=4, nhalo=5, dim=3, L=50.0)
    runners = [bf.BaryonifyShell(sh["catalog"], sh["field"], 20.0, sh["model"], verbose=False),
               bf.PaintProfilesShell(sh["catalog"], sh["field"], 20.0, sh["paint_model"], verbose=False),
               bf.BaryonifyGrid(gr["catalog"], gr["field"], 20.0, gr["model"], verbose=False),
               bf.PaintProfilesGrid(gr["catalog"], gr["field"], 20.0, gr["paint_model"], verbose=False),
               bf.BaryonifySnapshot(sn["catalog"], sn["field"], 5.0, sn["model"], verbose=False)]
    for r in runners:
        r2 = pickle.loads(pickle.dumps(r))
        assert type(r2) is type(r) and r2.epsilon_max == r.epsilon_max
        assert np.array_equal(r2.model.raw_input_r_range, r.model.raw_input_r_range)


def test_pin_in_place_state_machine(monkeypatch):
    """_device.pin_in_place without a GPU (the library calls are stubbed): first sight stages, second sight registers the
    range that is used, slices of a registered range need nothing, a failed or already page-locked range is remembered,
    and the registration / the sighting end with the array."""
    import gc
    import torch
    from baryonification_b200 import _device as dv

    class FakeLib:
        def __init__(self):
            self.pinned, self.rc = {}, 0

        def bf_host_pin(self, ptr, nbytes):
            if self.rc == 0:
                self.pinned[ptr.value] = nbytes
            return self.rc

        def bf_host_unpin(self, ptr):
            self.pinned.pop(ptr.value)
            return 0

    fake = FakeLib()
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(dv._lib, "load", lambda require_cuda=False: fake)
    monkeypatch.setattr(dv, "_PIN_MIN_BYTES", 1 << 10)
    monkeypatch.delenv("BF_PIN_INPUTS", raising=False)
    n0, s0 = len(dv._pin_done), len(dv._pin_seen)

    a = np.zeros(4096)
    assert not dv.pin_in_place(a[:10])                      # too small
    assert not dv.pin_in_place(a[::2])                      # not contiguous
    assert not dv.pin_in_place(a) and not fake.pinned       # first sight: staged
    assert not dv.pinned_already(a)
    assert dv.pin_in_place(a.reshape(64, 64))               # second sight (another view of the same range): registered
    assert fake.pinned == {a.ctypes.data: a.nbytes} and dv.pinned_already(a)
    assert dv.pin_in_place(a[1000:3000]) and len(fake.pinned) == 1   # a slice of a registered range
    del a
    gc.collect()
    assert not fake.pinned and len(dv._pin_done) == n0 and len(dv._pin_seen) == s0   # ended with the array

    b = np.zeros(4096)
    sl = b[512:2048]                                        # a rank's slice of a map: only that range is registered
    assert not dv.pin_in_place(sl) and dv.pin_in_place(b[512:2048])
    assert fake.pinned == {b.ctypes.data + 512 * 8: 1536 * 8}
    del b, sl
    gc.collect()
    assert not fake.pinned and len(dv._pin_seen) == s0

    c = np.zeros(4096)
    fake.rc = 1                                             # already page-locked by its owner: usable, nothing to undo
    assert not dv.pin_in_place(c) and dv.pin_in_place(c) and dv.pin_in_place(c) and dv.pinned_already(c)
    assert not fake.pinned
    d = np.zeros(4096)
    fake.rc = -1                                            # cannot be registered: remembered, staged from then on
    assert not dv.pin_in_place(d) and not dv.pin_in_place(d) and not dv.pin_in_place(d)
    fake.rc = 0
    assert not dv.pin_in_place(d) and not fake.pinned
    monkeypatch.setenv("BF_PIN_INPUTS", "1")                # register at first sight
    e = np.zeros(4096)
    assert dv.pin_in_place(e) and len(fake.pinned) == 1
    monkeypatch.setenv("BF_PIN_INPUTS", "0")                # off
    f = np.zeros(4096)
    assert not dv.pin_in_place(f) and not dv.pin_in_place(f)
    del c, d, e, f
    gc.collect()
    assert not fake.pinned and len(dv._pin_done) == n0 and len(dv._pin_seen) == s0
