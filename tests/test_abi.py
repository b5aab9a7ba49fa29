"""CPU: the C-ABI shared library loads without a GPU and exports every symbol include/*.h declares
(no compute calls here).  Also: the product refuses to run without CUDA instead of falling back."""
import ctypes as C
import glob
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = re.sub(r"/\*.*?\*/", " ", open(h).read(), flags=re.S)
        src = re.sub(r"//[^\n]*", " ", src)
        src = re.sub(r"typedef\s+struct\s*\{.*?\}\s*\w+\s*;", " ", src, flags=re.S)
        names |= set(re.findall(r"\b(bf_[A-Za-z0-9_]+)\s*\(", src))
    return sorted(names)


def test_header_declares_the_path():
    syms = declared_symbols()
    for must in ("bf_table_collapse_zM", "bf_healpix_query_disc", "bf_healpix_ang2pix", "bf_shell_bin_halos",
                 "bf_shell_gather", "bf_shell_regrid", "bf_grid_regrid", "bf_snap_displace", "bf_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    from baryonification_b200 import _lib, build
    build.build()
    lib = C.CDLL(_lib.SO_PATH)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing
    # and the ctypes binding table covers the header exactly
    assert sorted(_lib.EXPORTED) == declared_symbols()


def test_host_only_entry_points():
    """Entry points that take no device pointers work on a CPU-only box."""
    from baryonification_b200 import _lib
    lib = _lib.load()
    assert lib.bf_version() >= 100
    assert lib.bf_shell_num_bands(4096) == (4 * 4096 - 1 + 7) // 8
    f, l = C.c_int64(), C.c_int64()
    nb = lib.bf_shell_num_bands(64)
    assert lib.bf_shell_band_pixel_range(64, 0, nb, C.byref(f), C.byref(l)) == 0
    assert (f.value, l.value) == (0, 12 * 64 * 64)
    assert lib.bf_shell_band_pixel_range(64, 3, 2, C.byref(f), C.byref(l)) == -1  # BF_E_INVALID
    assert b"band range" in lib.bf_last_error()
    assert lib.bf_shell_bin_ws_bytes(64, 1000, 5000) > 0
    assert lib.bf_snap_grid_cells(3, 100.0, 7.0) == 14 and lib.bf_snap_grid_cells(2, 100.0, 1e-9) == 4096
    assert lib.bf_snap_ws_bytes(3, 1000, 100.0, 14, 5000) > 0 and lib.bf_snap_ws_bytes(3, 10, 100.0, 257, 0) == 0


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    import baryonification_b200 as bf
    from baryonification_b200 import _lib, synthetic as syn
    cfg = syn.config_shell(nside=8, nhalo=10)
    with pytest.raises(_lib.BFError, match="no CPU fallback"):
        bf.PaintProfilesShell(cfg["catalog"], cfg["field"], 20.0, cfg["paint_model"], verbose=False).process()
    with pytest.raises(RuntimeError, match="no CPU evaluation path"):
        cfg["model"].displacement(1.0, 1e14, 1.0)


def test_product_never_imports_the_oracle():
    for py in glob.glob(os.path.join(ROOT, "baryonification_b200", "**", "*.py"), recursive=True):
        src = open(py).read()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), py


def test_every_exported_entry_point_is_documented_in_integration_md():
    """INTEGRATION.md maps each C entry point to the reference lines it replaces (or says that it has no counterpart)."""
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "baryonification_b200.h")).read()
    doc = open(os.path.join(root, "INTEGRATION.md")).read()
    names = sorted(set(re.findall(r"\b(bf_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) > 50
    missing = [n for n in names if n not in doc]
    assert not missing, missing
