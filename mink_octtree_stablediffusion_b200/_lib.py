"""ctypes binding of libmink_b200.so (C-ABI: include/mink_b200.h).

ctypes releases the GIL around every foreign call, which mirrors the
`py::call_guard<py::gil_scoped_release>` on every reference op binding
(pybind/extern.hpp:517-646).  There is no CPU fallback: if the library is
missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "lib" / "libmink_b200.so"

MK_F32, MK_F64, MK_BF16, MK_F16 = 0, 1, 2, 3
MK_POOL_SUM, MK_POOL_AVG, MK_POOL_MAX = 0, 1, 2
MK_BCAST_ADD, MK_BCAST_MUL = 0, 1
MK_INSERT_UNIQUE, MK_INSERT_INT64_MAPS = 1, 2

_p = C.c_void_p
_i64 = C.c_int64
_i = C.c_int

# name -> (restype, argtypes); must list every symbol include/mink_b200.h declares
PROTOTYPES = {
    "mk_last_error": (C.c_char_p, []),
    "mk_version": (_i, []),
    "mk_launch_count": (_i64, []),
    "mk_has_tcgen05": (_i, []),
    "mk_hash_capacity": (_i64, [_i64]),
    "mk_insert_workspace_bytes": (_i64, [_i64]),
    "mk_map_insert": (_i, [_p, _i64, _i, _p, _i64, _p, _p, _p, _p, _p, _p, _p]),
    "mk_insert2_workspace_bytes": (_i64, [_i64]),
    "mk_map_insert2": (_i, [_p, _i64, _i, _p, _i64, _i, _p, _p, _p, _p, _p, _p, _p]),
    "mk_map_lookup": (_i, [_p, _i64, _p, _i64, _i, _p, _p]),
    "mk_coords_stride": (_i, [_p, _i64, _i, _p, _p, _p]),
    "mk_coords_expand": (_i, [_p, _i64, _i, _p, _i, _p, _p]),
    "mk_coords_origin": (_i, [_p, _i64, _i, _p, _p]),
    "mk_select_workspace_bytes": (_i64, [_i64]),
    "mk_select_flagged": (_i, [_p, _i64, _p, _p, _p, _p]),
    "mk_select_static_max_rows": (_i64, []),
    "mk_select_flagged_static": (_i, [_p, _i64, _p, _p, _p]),
    "mk_rows_gather": (_i, [_p, _p, _i64, _i64, _p, _p]),
    "mk_rows_scatter": (_i, [_p, _p, _i64, _i64, _p, _p]),
    "mk_kmap_build": (_i, [_p, _i64, _i, _p, _i64, _p, _i, _p, _i64, _p, _p]),
    "mk_kmap_invert": (_i, [_p, _i64, _i64, _i, _p, _i64, _i64, _p]),
    "mk_kmap_children": (_i, [_i64, _i, _i64, _i64, _p, _p, _p, _p, _p]),
    "mk_kmap_pairs_workspace_bytes": (_i64, [_i64, _i]),
    "mk_kmap_pairs": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _p, _p]),
    "mk_conv_workspace_bytes": (_i64, [_i, _i64, _i, _i, _i]),
    "mk_kmap_tile_mask_words": (_i64, [_i]),
    "mk_kmap_tile_mask": (_i, [_p, _i64, _i64, _i, _p, _p]),
    "mk_conv_fwd": (_i, [_i, _p, _p, _i, _i, _p, _p, _i64, _i64, _i64, _i, _i, _i, _p, _p, _p]),
    "mk_conv_fwd_rows": (_i, [_i, _p, _p, _i, _i, _p, _p, _i64, _i64, _i64, _i, _i, _i, _p, _p, _p, _p]),
    "mk_conv_tc_eligible": (_i, [_i, _i64, _i, _i, _i]),
    "mk_kmap_tile_order_workspace_bytes": (_i64, [_i64]),
    "mk_kmap_tile_order": (_i, [_p, _i64, _i64, _i, _p, _i, _i, _p, _p, _p, _p]),
    "mk_conv_wgrad": (_i, [_i, _p, _p, _p, _p, _i64, _i64, _i64, _i, _i, _i, _p, _p, _p]),
    "mk_conv_wgrad_rows": (_i, [_i, _p, _p, _p, _p, _p, _i64, _i64, _i64, _i, _i, _i, _p, _p, _p]),
    "mk_conv_workspace_bytes2": (_i64, [_i, _i64, _i64, _i, _i, _i]),
    "mk_conv_fwd_ex": (_i, [_i, _p, _p, _i, _i, _p, _p, _i64, _i64, _i64, _i, _i, _i, _p, _p, _p, _i64, _p]),
    "mk_conv_wgrad_ex": (_i, [_i, _p, _p, _p, _p, _p, _i64, _i64, _i64, _i, _i, _i, _p, _i, _p, _i64, _p]),
    "mk_conv_rows_eligible": (_i, [_i, _i64, _i, _i, _i]),
    "mk_conv_last_path": (_i, []),
    "mk_conv_force_simt": (_i, [_i]),
    "mk_pool_fwd": (_i, [_i, _i, _p, _p, _i64, _i64, _i, _i, _p, _p, _p, _p]),
    "mk_pool_bwd": (_i, [_i, _i, _p, _p, _p, _i64, _i64, _i, _i, _p, _i64, _p, _p]),
    "mk_segment_workspace_bytes": (_i64, [_i, _i, _i64]),
    "mk_global_pool_fwd": (_i, [_i, _i, _p, _p, _i64, _i, _i64, _p, _p, _p, _p, _p]),
    "mk_global_pool_bwd": (_i, [_i, _i, _p, _p, _p, _i64, _i, _i64, _p, _p, _p]),
    "mk_broadcast_fwd": (_i, [_i, _i, _p, _p, _p, _i64, _i, _p, _p]),
    "mk_broadcast_bwd": (_i, [_i, _i, _p, _p, _p, _p, _i64, _i, _i64, _p, _p, _p, _p]),
    "mk_seg_sums": (_i, [_i, _p, _p, _p, _p, _p, _i64, _i, _i64, _p, _p]),
    "mk_bn_coef_fwd": (_i, [_i, _p, _p, _p, C.c_double, _p, _p, C.c_double, _i, _p, _p, _p, _p, _p, _p, C.c_double, _p]),
    "mk_bn_coef_bwd": (_i, [_i, _p, _p, _p, _p, C.c_double, _p, _i, _p, _p, _p, _p, _p, _p]),
    "mk_seg_affine": (_i, [_i, _p, _p, _p, _i64, _i, _p, _p, _p, _p, _p]),
    "mk_gn_coef_fwd": (_i, [_i, _p, _p, _p, _p, _p, C.c_double, _i, _i, _i, _i64, _p, _p, _p, _p, _p, _p]),
    "mk_gn_coef_bwd": (_i, [_i, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i64, _p, _p, _p, _p, _p, _p]),
}

_lib = None


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load (building in-tree first if needed) the C-ABI library."""
    global _lib
    if _lib is not None:
        return _lib
    override = os.environ.get("MK_B200_LIB")      # developer aid: A/B a differently built library
    if override:
        lib = C.CDLL(override)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
        return lib
    if not LIB_PATH.exists():
        if not build_if_missing:
            raise RuntimeError(f"{LIB_PATH} is missing; run `python -m mink_octtree_stablediffusion_b200.build`")
        from . import build as _build
        _build.build_library()
    else:
        # a library older than its sources must not be measured or tested silently: the stamp beside it
        # names the source digests it was linked from
        from . import build as _build
        stale = _build.stale_reason()
        if stale:
            if build_if_missing and _build.have_nvcc():
                _build.build_library()
            else:
                raise RuntimeError(f"{LIB_PATH} is stale ({stale}); run `python -m mink_octtree_stablediffusion_b200.build`")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)          # AttributeError => symbol missing: fail loudly
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().mk_last_error()
        raise RuntimeError(f"mink_b200 error {rc}: {msg.decode() if msg else '?'}")
