"""GPU: the allocation-free entry point ``bf_pk_bin_static`` (scratch in a file-scope device array of 8 MiB, no caller
workspace, no counterpart in the reference) against the existing
``bf_pk_bin`` with a caller workspace -- through the C ABI, on two calls in a row with different inputs (the array keeps
the first call's contents), and the limits it must reject.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _spectrum(ngrd, seed):
    import torch
    g = torch.Generator(device="cuda").manual_seed(seed)
    grid = torch.rand((ngrd,) * 3, dtype=torch.float64, device="cuda", generator=g)
    return torch.fft.rfftn(grid)  # complex128, a permuted dense view: both entry points read it through its strides


def _consts(ngrd, nk, L=100.0):
    kbins = np.linspace(2 * np.pi / L, 2 * np.pi / L * ngrd / 2, nk + 1)
    d = 1 / (2 * np.pi / L) / ngrd
    return 1.0 / (ngrd * d), float(kbins[0]), float(kbins[1] - kbins[0])


def test_pk_bin_static_equals_pk_bin_on_two_calls_in_a_row(gpu):
    import torch
    from baryonification_b200 import _device as dv, _lib
    lib = _lib.load(require_cuda=True)
    assert lib.bf_pk_bin_static_max_bytes() == 8 << 20
    for call, (ngrd, nk, seed) in enumerate([(96, 180, 1), (64, 256, 2), (96, 180, 1)]):  # 2nd: other shape, the most nk
        F = _spectrum(ngrd, seed)
        val, k0, dk = _consts(ngrd, nk)
        strides = (C.c_int64 * 3)(*F.stride())
        # everything the two calls need exists BEFORE them: the static entry point is handed inputs and outputs only
        nb = lib.bf_pk_bin_ws_bytes(ngrd, nk)
        ws = torch.empty(nb, dtype=torch.uint8, device="cuda")
        want, got = dv.empty((3, nk)), dv.empty((3, nk))
        got.fill_(float("nan"))
        _lib.call("bf_pk_bin", ngrd, dv.ptr(F), strides, val, k0, dk, nk, dv.ptr(want), dv.ptr(ws), nb, dv.stream_ptr())
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        _lib.call("bf_pk_bin_static", ngrd, dv.ptr(F), strides, val, k0, dk, nk, dv.ptr(got), dv.stream_ptr())
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        w, g = want.cpu().numpy(), got.cpu().numpy()
        assert np.array_equal(g[2], w[2]) and w[2].sum() > 0.9 * (ngrd ** 3) * 0.5, call  # mode counts: exact
        assert np.allclose(g, w, rtol=1e-13, atol=0), call                                   # the tests' tolerance ...
        assert np.array_equal(g, w), call                                                    # ... and in fact the same bits


def test_pk_bin_static_rejects_what_does_not_fit(gpu):
    from baryonification_b200 import _device as dv, _lib
    lib = _lib.load(require_cuda=True)
    F = _spectrum(32, 3)
    strides = (C.c_int64 * 3)(*F.stride())
    out = dv.empty((3, 512))
    val, k0, dk = _consts(32, 300)
    rc = lib.bf_pk_bin_static(32, dv.ptr(F), strides, val, k0, dk, 300, dv.ptr(out), dv.stream_ptr())
    assert rc == -1 and b"outside [1, 256]" in lib.bf_last_error()
    rc = lib.bf_pk_bin_static(32, dv.ptr(F), strides, val, k0, dk, 0, dv.ptr(out), dv.stream_ptr())
    assert rc == -1
