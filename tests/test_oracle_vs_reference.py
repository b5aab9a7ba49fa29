"""Build-container only (/root/reference present): the oracle against the reference package executed verbatim
(oracle/ref_shim.py supplies pyccl / healpy), on more and larger seeded cases than the committed fixtures."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.reference


@pytest.fixture(scope="module")
def bfg():
    from oracle import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("reference tree not present (GPU box)")
    return ref_shim.import_reference()


def _mg():
    import importlib.util, os
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(cases.GOLDEN, "make_golden.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


LIVE = [
    ("shell_baryonify", dict(nside=128, nhalo=500, epsilon_max=10.0, seed=11)),
    ("shell_baryonify", dict(nside=4, nhalo=100, epsilon_max=40.0, seed=12)),  # discs swallow the poles
    ("shell_paint", dict(nside=128, nhalo=500, epsilon_max=20.0, zero_map=True, zero_beyond=50.0, seed=13)),
    ("grid_baryonify", dict(npix=128, nhalo=300, dim=2, epsilon_max=20.0, seed=14)),
    ("grid_paint", dict(npix=128, nhalo=300, dim=2, epsilon_max=20.0, seed=15)),
    ("grid_baryonify", dict(npix=32, nhalo=60, dim=3, epsilon_max=6.0, L=120.0, seed=16)),
    ("grid_paint", dict(npix=32, nhalo=60, dim=3, epsilon_max=6.0, L=120.0, seed=17)),
    ("grid_baryonify", dict(npix=128, nhalo=300, dim=2, epsilon_max=20.0, seed=24, ellipticity=True)),
    ("grid_paint", dict(npix=128, nhalo=300, dim=2, epsilon_max=20.0, seed=25, ellipticity=True)),
    ("shell_paint", dict(nside=64, nhalo=400, epsilon_max=20.0, zero_map=True, zero_beyond=50.0, seed=26, p_keys=True)),
    ("grid_paint", dict(npix=96, nhalo=250, dim=2, epsilon_max=20.0, seed=27, p_keys=True)),
    ("grid_paint", dict(npix=24, nhalo=60, dim=3, epsilon_max=6.0, L=120.0, seed=28, p_keys=True)),
    ("shell_baryonify", dict(nside=64, nhalo=300, epsilon_max=20.0, seed=31, rdelta=True)),  # Rdelta_sampling tables
    ("grid_baryonify", dict(npix=96, nhalo=200, dim=2, epsilon_max=20.0, seed=32, rdelta=True)),
    ("snapshot", dict(npart_side=20, nhalo=120, dim=3, epsilon_max=5.0, L=100.0, seed=33, rdelta=True)),
    ("snapshot", dict(npart_side=24, nhalo=150, dim=3, epsilon_max=5.0, L=100.0, seed=18)),
    ("snapshot", dict(npart_side=100, nhalo=100, dim=2, epsilon_max=5.0, L=300.0, seed=19)),
]


@pytest.mark.parametrize("kind,kw", LIVE, ids=[f"{k}-{i}" for i, (k, _) in enumerate(LIVE)])
def test_oracle_equals_live_reference(bfg, kind, kw):
    from oracle import runners as orc
    mg = _mg()
    cfg = mg.build_case(kind, kw)
    with np.errstate(all="ignore"):
        want = mg.run_reference(bfg, kind, cfg)
    _, got = cases.run(orc, kind, cfg)
    cases.check(kind, got, want, cfg)


def test_reference_healpix_regrid_kernel(bfg):
    """regrid_pixels_hpix (numba, HealpixRunner.py:13-67) == oracle bfo_regrid_shell on random offsets."""
    from oracle import healpix as ohp, runners as orc
    from BaryonForge.Runners.HealpixRunner import regrid_pixels_hpix
    nside, rng = 16, np.random.default_rng(5)
    npix = 12 * nside * nside
    orig = rng.lognormal(size=npix)
    orig[rng.random(npix) < 0.2] = 0.0
    off = rng.normal(scale=0.05, size=(npix, 3))
    vec = np.stack(ohp.pix2vec(nside, np.arange(npix)), 1) + off
    lon, lat = ohp.vec2ang(vec, lonlat=True)
    p = np.flatnonzero(orig > 0)
    cp, cw = ohp.get_interp_weights(nside, lon[p], lat[p], lonlat=True)
    want = regrid_pixels_hpix(np.zeros(npix), orig[p], cp.T, cw.T)
    got = np.zeros(npix)
    orc._lib().bfo_regrid_shell(nside, orig, np.ascontiguousarray(off), got)
    assert np.allclose(got, want, rtol=1e-13, atol=0)


def test_make_map_host_form_equals_the_reference(bfg):
    """utils/io.py:657-670 -- the repo's ParticleSnapshot.make_map(device="host") is the reference's numpy call."""
    from baryonification_b200 import synthetic as syn
    rng = np.random.default_rng(5)
    for dim in (2, 3):
        xyz = [rng.uniform(0, 100.0, 20_000) for _ in range(dim)]
        xyz[0][:5] = [0.0, 100.0, 50.0, 25.0, 100.0]
        M = rng.lognormal(0, 1, 20_000)
        kw = dict(x=xyz[0], y=xyz[1], z=xyz[2] if dim == 3 else None, M=M, L=100.0, redshift=0.3, cosmo=syn.COSMO)
        want = bfg.utils.io.ParticleSnapshot(**kw).make_map(16)
        got = syn.ParticleSnapshot(**kw).make_map(16, device="host")
        assert np.array_equal(got, want)
