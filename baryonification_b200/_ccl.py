"""Resolve the cosmology back-end used for the per-halo *input* scalars.

The reference calls pyccl for R_Delta and D_A only (see ``_miniccl``).  If the
real ``pyccl`` imports, it is used, so a BaryonForge user gets bit-identical
halo radii.  Otherwise the documented flat-wCDM stand-in is used.
"""
from __future__ import annotations

import sys

_backend = None


def get_ccl():
    global _backend
    if _backend is None:
        try:
            import pyccl as ccl  # real library, or a stand-in a test harness registered
            _backend = ccl
        except Exception:
            from . import _miniccl
            mods = _miniccl.as_module("pyccl")
            # do NOT register in sys.modules: the product must not shadow a later real install
            _backend = mods["pyccl"]
    return _backend


def default_mass_def():
    return get_ccl().halos.massdef.MassDef(200, "critical")


def backend_name():
    ccl = get_ccl()
    return "miniccl" if getattr(ccl, "__miniccl__", False) else "pyccl-" + getattr(ccl, "__version__", "?")
