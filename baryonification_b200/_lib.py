"""ctypes binding of the C ABI in include/baryonification_b200.h.

There is no CPU fallback: if the CUDA extension is missing or CUDA is unavailable the runners
raise.  Device memory, streams and multi-GPU plumbing come from PyTorch; every compute call goes
through the `extern "C"` entry points with raw device pointers.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# BF_LIB_PATH: a differently-compiled build of the same library (kernel tuning experiments, scripts/build_variant.py)
SO_PATH = os.environ.get("BF_LIB_PATH") or os.path.join(HERE, "libbaryonification_b200.so")

_dbl_p = C.c_void_p  # raw device pointers
_i64 = C.c_int64
_lib = None


class BFError(RuntimeError):
    pass


class Table(C.Structure):
    _fields_ = [("gz", C.c_void_p), ("gM", C.c_void_p), ("gr", C.c_void_p), ("values", C.c_void_p),
                ("nz", _i64), ("nM", _i64), ("nr", _i64), ("lnr_first", C.c_double), ("lnr_last", C.c_double)]


MAX_AXES = 6


class TableND(C.Structure):
    _fields_ = [("naxes", C.c_int), ("grid", C.c_void_p * MAX_AXES), ("n", _i64 * MAX_AXES), ("values", C.c_void_p),
                ("nr", _i64)]


class ShellHalos(C.Structure):
    _fields_ = [("vec", C.c_void_p), ("lon", C.c_void_p), ("lat", C.c_void_p), ("radius", C.c_void_p),
                ("D", C.c_void_p), ("a", C.c_void_p), ("cut", C.c_void_p), ("lnr_shift", C.c_void_p),
                ("rows", C.c_void_p), ("order", C.c_void_p), ("nh", _i64)]


class GridHalos(C.Structure):
    _fields_ = [("pos", C.c_void_p), ("Nsize", C.c_void_p), ("cut", C.c_void_p), ("lnr_shift", C.c_void_p),
                ("rows", C.c_void_p), ("Rmat", C.c_void_p), ("nh", _i64)]


class SnapHalos(C.Structure):
    _fields_ = [("pos", C.c_void_p), ("Rq", C.c_void_p), ("cut", C.c_void_p), ("lnr_shift", C.c_void_p),
                ("rows", C.c_void_p), ("nh", _i64)]


_P = C.POINTER
_SIGS = {
    "bf_last_error": (C.c_char_p, []),
    "bf_version": (C.c_int, []),
    "bf_launch_count": (C.c_ulonglong, []),
    "bf_probe_fp64": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_ulonglong), C.c_void_p]),
    "bf_device_sm_count": (C.c_int, [_P(C.c_int)]),
    "bf_host_pin": (C.c_int, [C.c_void_p, C.c_size_t]),
    "bf_host_unpin": (C.c_int, [C.c_void_p]),
    "bf_table_collapse_zM": (C.c_int, [_P(Table), _i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_table_collapse_nd": (C.c_int, [_P(TableND), _i64, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_table_readout": (C.c_int, [_P(Table), _dbl_p, _i64, _dbl_p, C.c_double, C.c_double, C.c_int, _dbl_p,
                                   C.c_void_p]),
    "bf_healpix_ang2pix": (C.c_int, [_i64, _i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_healpix_pix2ang": (C.c_int, [_i64, _i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_healpix_pix2vec": (C.c_int, [_i64, _i64, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_healpix_ang2vec_lonlat": (C.c_int, [_i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_healpix_interp_weights_lonlat": (C.c_int, [_i64, _i64, _dbl_p, _dbl_p, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_healpix_query_disc": (C.c_int, [_i64, _i64, _dbl_p, _dbl_p, _dbl_p, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_shell_paint_atomic": (C.c_int, [_i64, _P(ShellHalos), _P(Table), _dbl_p, _dbl_p, C.c_void_p]),
    "bf_shell_baryonify_atomic": (C.c_int, [_i64, _P(ShellHalos), _P(Table), _dbl_p, _dbl_p, C.c_void_p]),
    "bf_shell_num_bands": (_i64, [_i64]),
    "bf_shell_band_pixel_range": (C.c_int, [_i64, _i64, _i64, _P(_i64), _P(_i64)]),
    "bf_shell_bin_ws_bytes": (C.c_size_t, [_i64, _i64, _i64]),
    "bf_shell_bin_halos": (C.c_int, [_i64, C.c_int, _P(ShellHalos), _i64, _i64, _i64, C.c_void_p, C.c_size_t,
                                     _P(_i64), C.c_void_p]),
    "bf_shell_bin_begin": (C.c_int, [_i64, C.c_int, _P(ShellHalos), _i64, _i64, _i64, C.c_void_p, C.c_size_t, C.c_void_p]),
    "bf_shell_bin_add": (C.c_int, [_i64, C.c_int, _P(ShellHalos), _i64, _i64, _i64, _i64, _i64, C.c_void_p, C.c_size_t,
                                   C.c_void_p]),
    "bf_shell_bin_end": (C.c_int, [_i64, C.c_int, _P(ShellHalos), _i64, _i64, _i64, C.c_void_p, C.c_size_t, _P(_i64),
                                   C.c_void_p]),
    "bf_shell_gather": (C.c_int, [_i64, C.c_int, _P(ShellHalos), _P(Table), _i64, _i64, C.c_void_p, _dbl_p, _i64,
                                  C.c_int, _dbl_p, C.c_void_p]),
    "bf_shell_regrid": (C.c_int, [_i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_shell_regrid_range": (C.c_int, [_i64, _i64, _i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_shell_regrid_sorted_ws_bytes": (C.c_size_t, [_i64, _i64]),
    "bf_shell_regrid_sorted": (C.c_int, [_i64, _i64, _i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "bf_shell_regrid_gather_ws_bytes": (C.c_size_t, [_i64]),
    "bf_shell_regrid_gather": (C.c_int, [_i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_int),
                                         C.c_void_p]),
    "bf_shell_regrid_gather_emit": (C.c_int, [_i64, _dbl_p, _dbl_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_int),
                                              C.c_void_p]),
    "bf_shell_regrid_gather_emit_range": (C.c_int, [_i64, _i64, _i64, _dbl_p, _dbl_p, C.c_void_p, C.c_size_t,
                                                    C.POINTER(C.c_int), C.c_void_p]),
    "bf_shell_regrid_gather_emit_part": (C.c_int, [_i64, _i64, _i64, _dbl_p, _dbl_p, C.c_void_p, C.c_size_t,
                                                    C.POINTER(C.c_int), C.c_void_p]),
    "bf_shell_regrid_gather_range": (C.c_int, [_i64, C.c_void_p, C.c_size_t, C.POINTER(C.c_int), _i64, _i64, _dbl_p,
                                               C.c_void_p]),
    "bf_grid_baryonify_atomic": (C.c_int, [C.c_int, _i64, _dbl_p, C.c_double, _P(GridHalos), _P(Table), _dbl_p,
                                           _dbl_p, C.c_void_p]),
    "bf_grid_paint_atomic": (C.c_int, [C.c_int, _i64, _dbl_p, C.c_double, _P(GridHalos), _P(Table), _dbl_p,
                                       _dbl_p, C.c_void_p]),
    "bf_grid_bin_ws_bytes": (C.c_size_t, [C.c_int, _i64, _i64, _i64]),
    "bf_grid_bin_halos": (C.c_int, [C.c_int, _i64, _dbl_p, _P(GridHalos), _i64, _i64, C.c_void_p, C.c_size_t,
                                    _P(_i64), C.c_void_p]),
    "bf_grid_gather": (C.c_int, [C.c_int, C.c_int, _i64, C.c_double, _P(GridHalos), _P(Table), _i64, C.c_void_p,
                                 _dbl_p, _dbl_p, C.c_void_p]),
    "bf_grid_gather_rows": (C.c_int, [C.c_int, C.c_int, _i64, C.c_double, _P(GridHalos), _P(Table), _i64, C.c_void_p,
                                      _dbl_p, _i64, _i64, _dbl_p, C.c_void_p]),
    "bf_grid_regrid_window": (C.c_int, [C.c_int, _i64, _dbl_p, _i64, _i64, C.c_void_p, C.c_size_t, C.POINTER(C.c_int),
                                        C.c_void_p]),
    "bf_grid_regrid_gather_rows": (C.c_int, [C.c_int, _i64, _dbl_p, _dbl_p, _dbl_p, C.POINTER(C.c_int), _i64, _i64,
                                             C.POINTER(C.c_int), C.c_void_p]),
    "bf_grid_regrid": (C.c_int, [C.c_int, _i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_grid_regrid_sorted_ws_bytes": (C.c_size_t, [C.c_int, _i64]),
    "bf_grid_regrid_sorted": (C.c_int, [C.c_int, _i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "bf_grid_regrid_gather_ws_bytes": (C.c_size_t, []),
    "bf_grid_regrid_gather": (C.c_int, [C.c_int, _i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p, C.c_size_t,
                                        C.POINTER(C.c_int), C.c_void_p]),
    "bf_snap_grid_cells": (_i64, [C.c_int, C.c_double, C.c_double]),
    "bf_snap_ws_bytes": (C.c_size_t, [C.c_int, _i64, C.c_double, _i64, _i64]),
    "bf_snap_bin_halos": (C.c_int, [C.c_int, C.c_double, _i64, _P(SnapHalos), _i64, C.c_void_p, C.c_size_t, _P(_i64),
                                    C.c_void_p]),
    "bf_snap_displace": (C.c_int, [C.c_int, _i64, _dbl_p, _dbl_p, _dbl_p, C.c_double, C.c_double, _P(SnapHalos),
                                   _P(Table), _i64, C.c_void_p, _dbl_p, _dbl_p, _dbl_p, _dbl_p, C.c_void_p]),
    "bf_snap_pair_list": (C.c_int, [C.c_int, _i64, _dbl_p, _dbl_p, _dbl_p, C.c_double, C.c_double, _P(SnapHalos),
                                    _P(Table), _i64, C.c_void_p, _dbl_p, _dbl_p, _dbl_p, C.c_void_p, _i64, C.c_void_p,
                                    C.c_void_p]),
    "bf_snap_displace_records": (C.c_int, [C.c_int, _i64, C.c_void_p, C.c_double, C.c_double, _P(SnapHalos), _P(Table),
                                           _i64, C.c_void_p, C.c_void_p, _dbl_p, C.c_void_p]),
    "bf_make_map_ws_bytes": (C.c_size_t, [_i64]),
    "bf_make_map": (C.c_int, [C.c_int, _i64, _dbl_p, _dbl_p, _dbl_p, _dbl_p, C.c_void_p, _dbl_p, _i64, _dbl_p, C.c_void_p,
                              C.c_size_t, C.c_void_p]),
    "bf_pk_ngp_deposit": (C.c_int, [_i64, _dbl_p, _dbl_p, _dbl_p, C.c_void_p, C.c_double, _i64, _dbl_p, C.c_void_p]),
    "bf_pk_bin_ws_bytes": (C.c_size_t, [_i64, _i64]),
    "bf_pk_bin": (C.c_int, [_i64, C.c_void_p, _P(_i64), C.c_double, C.c_double, C.c_double, _i64, _dbl_p, C.c_void_p,
                            C.c_size_t, C.c_void_p]),
    "bf_pk_bin_static_max_bytes": (C.c_size_t, []),
    "bf_pk_bin_static": (C.c_int, [_i64, C.c_void_p, _P(_i64), C.c_double, C.c_double, C.c_double, _i64, _dbl_p,
                                   C.c_void_p]),
    "bf_pk_bin_part_ws_bytes": (C.c_size_t, [_P(_i64), _i64]),
    "bf_pk_bin_part": (C.c_int, [_i64, C.c_void_p, _P(_i64), _P(_i64), _P(_i64), C.c_double, C.c_double, C.c_double,
                                 _i64, _dbl_p, C.c_void_p, C.c_size_t, C.c_void_p]),
}

EXPORTED = tuple(_SIGS)


def load(require_cuda: bool = False):
    """dlopen the extension (needs no GPU).  Raises BFError when the .so is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise BFError(f"CUDA extension not built: {SO_PATH} is missing. Run `python -m baryonification_b200.build` "
                          "(or __graft_entry__.build()). There is no CPU fallback.")
        lib = C.CDLL(SO_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(lib, name)  # AttributeError here = header / library mismatch
            fn.restype, fn.argtypes = res, args
        _lib = lib
    if require_cuda:
        import torch
        if not torch.cuda.is_available():
            raise BFError("baryonification_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return _lib


def check(status: int, what: str = ""):
    if status != 0:
        msg = load().bf_last_error().decode(errors="replace")
        raise BFError(f"{what or 'bf call'} failed with status {status}: {msg}")


def call(name, *args):
    check(getattr(load(require_cuda=True), name)(*args), name)
