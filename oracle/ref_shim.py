"""oracle/ref_shim.py -- TEST INFRASTRUCTURE ONLY.

Lets the UNMODIFIED reference package (``/root/reference/BaryonForge``) import
and run in this container, where its two third-party dependencies are missing:

* ``pyccl``  -> ``baryonification_b200._miniccl`` (flat wCDM closed forms; the
  per-halo scalars R_Delta and D_A are *inputs* to both the oracle and the GPU
  path, so both sides see identical numbers);
* ``healpy`` -> ``oracle.healpix`` (C restatement of the HEALPix RING routines;
  parity unpinned against real healpy).

If the real modules import, they are left alone and the reference runs against
them (recorded by :func:`describe`).  ``/root/reference`` exists only in the
build container; nothing that runs on the GPU box may call
:func:`import_reference` (tests that do are skipped when the path is absent).
"""
from __future__ import annotations

import importlib
import os
import sys

REFERENCE_ROOT = os.environ.get("BF_REFERENCE_ROOT", "/root/reference")
_state = {"pyccl": None, "healpy": None}


def install():
    """Make ``import pyccl`` / ``import healpy`` succeed (real module or stand-in)."""
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(here)
    if root not in sys.path:
        sys.path.insert(0, root)
    if _state["pyccl"] is None:
        try:
            import pyccl  # noqa: F401
            _state["pyccl"] = "pyccl-" + getattr(pyccl, "__version__", "?")
        except Exception:
            from baryonification_b200 import _miniccl
            sys.modules.update(_miniccl.as_module("pyccl"))
            _state["pyccl"] = "miniccl"
    if _state["healpy"] is None:
        try:
            import healpy  # noqa: F401
            _state["healpy"] = "healpy-" + getattr(healpy, "__version__", "?")
        except Exception:
            from oracle import healpix as _hpx
            sys.modules["healpy"] = _hpx.as_module("healpy")
            _state["healpy"] = "restated"
    return dict(_state)


def describe():
    return dict(_state)


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "BaryonForge"))


def import_reference():
    """Import the reference package verbatim (build container only)."""
    if not reference_available():
        raise RuntimeError(f"reference tree not present at {REFERENCE_ROOT}")
    install()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return importlib.import_module("BaryonForge")
