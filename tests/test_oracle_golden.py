"""CPU: the oracle (oracle/*.c + oracle/runners.py) against the committed outputs of the reference itself
(tests/golden/*.npz, produced by tests/golden/make_golden.py running /root/reference verbatim)."""
import numpy as np
import pytest

import cases


@pytest.mark.parametrize("name", cases.case_names())
def test_oracle_matches_reference_fixture(name):
    from oracle import runners as orc
    kind, cfg = cases.build(name)
    _, got = cases.run(orc, kind, cfg)
    cases.check(kind, got, cases.golden(name), cfg)


@pytest.mark.parametrize("name", ["shell_paint_ns64", "grid2d_baryonify_64", "snapshot3d_16"])
def test_oracle_threaded_equals_serial(name):
    """The OpenMP leg used as bench.py's cpu_baseline gives the serial answer (atomics only reorder sums)."""
    from oracle import runners as orc
    kind, cfg = cases.build(name)
    r1, a = cases.run(orc, kind, cfg, nthreads=1)
    r4, b = cases.run(orc, kind, cfg, nthreads=4)
    assert r1.npairs == r4.npairs
    for k in a:
        assert np.allclose(a[k], b[k], rtol=1e-12, atol=1e-300)


def test_readout_matches_reference_fixture():
    """bfo_readout vs the reference's scipy RegularGridInterpolator readouts (BaryonCorrection.py:356-390,
    Tabulate.py:272-294): same NaN pattern (out of table, 0 * -inf), values to 1e-13."""
    from baryonification_b200 import synthetic as syn
    from oracle import runners as orc
    g = cases.golden("readout")
    dm = syn.displacement_model(0.05, 1.2, 20.0)
    pm = syn.profile_model(0.05, 1.2, zero_beyond=50.0)
    with np.errstate(all="ignore"):
        for i, (M, z) in enumerate(zip(g["M"], g["z"])):
            a = 1 / (1 + z)
            for got, want in ((orc.readout_displacement(dm, g["r"], M, a), g["displacement"][i]),
                              (orc.readout_profile(pm, g["r"], M, a, "2D"), g["projected"][i]),
                              (orc.readout_profile(pm, g["r"], M, a, "3D"), g["real"][i])):
                assert np.array_equal(np.isnan(got), np.isnan(want))
                ok = ~np.isnan(want)
                assert np.allclose(got[ok], want[ok], rtol=1e-13, atol=1e-300)
            assert np.isnan(g["displacement"][i]).any() or i != 2  # case 2 (M below the table) is all-NaN
