"""Pruning, pooling and broadcast `*GPU` ops of the `_C` surface
(pybind/extern.hpp:215-428) on the C-ABI library."""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from .backend import (BroadcastMode, CoordinateMapKey, PoolingMode, _POOL_KIND, _dtype_code, _ptr, _require,
                      _stream)
from .backend_ops import _inverse_table, _kernel_map, _resolve_out_key


# ---------------------------------------------------------------------------
# pruning (pybind/extern.hpp:411-428; MinkowskiPruning.py:38-72)
# ---------------------------------------------------------------------------
def PruningForwardGPU(in_feat, keep, in_key, out_key, mgr):
    """New map with the kept coordinates (row order preserved) + their features."""
    lib = mgr._lib
    m = mgr._get(in_key)
    _require(in_feat.is_cuda and in_feat.dim() == 2, "in_feat must be a 2-D CUDA tensor")
    _require(in_feat.size(0) == m.n, "in_feat.size(0) == p_map_manager->size(in_key)")
    _require(keep.dtype in (torch.bool, torch.uint8), "keep.dtype() must be kBool or kByte")
    _require(keep.numel() == m.n, "keep.numel() == in_feat.size(0)")
    in_feat = in_feat.contiguous()
    dev = in_feat.device
    keep_u8 = keep.to(device=dev).contiguous().view(-1)
    keep_u8 = keep_u8.view(torch.uint8) if keep_u8.dtype == torch.bool else keep_u8
    kept = torch.empty(max(m.n, 1), dtype=torch.int32, device=dev)
    cnt = torch.empty(1, dtype=torch.int64, device=dev)
    ws = torch.empty(int(lib.mk_select_workspace_bytes(m.n)), dtype=torch.uint8, device=dev)
    _lib.check(lib.mk_select_flagged(_ptr(keep_u8), m.n, _ptr(kept), _ptr(ws), ctypes.c_void_p(cnt.data_ptr()),
                                     _stream()))
    n_kept = int(cnt.item())
    kept = kept[:n_kept]
    coords = torch.empty((n_kept, m.coords.shape[1]), dtype=torch.int32, device=dev)
    out = torch.empty((n_kept, in_feat.size(1)), dtype=in_feat.dtype, device=dev)
    if n_kept > 0:
        _lib.check(lib.mk_rows_gather(_ptr(m.coords), _ptr(kept), n_kept, m.coords.shape[1] * 4, _ptr(coords),
                                      _stream()))
        _lib.check(lib.mk_rows_gather(_ptr(in_feat), _ptr(kept), n_kept, in_feat.size(1) * in_feat.element_size(),
                                      _ptr(out), _stream()))
    if not out_key.is_key_set():
        sid = in_key.get_key()[1]
        key, _, _ = mgr._new_map_from_coords(coords, m.tensor_stride, sid + "p" if sid else "p", unique=True)
        out_key.set_key(*key.get_key())
        mgr._get(out_key).on_grid = m.on_grid
    else:
        _require(mgr.size(out_key) == n_kept, "pruning: preset out_key has a different size")
    if not hasattr(mgr, "_prune_rows"):
        mgr._prune_rows = {}
    mgr._prune_rows[(in_key._tuple(), out_key._tuple())] = kept
    return out


def PruningBackwardGPU(grad_out_feat, in_key, out_key, mgr):
    """zeros[N_in, C] with the kept rows copied back (MinkowskiPruning.py:62-72)."""
    lib = mgr._lib
    kept = getattr(mgr, "_prune_rows", {}).get((in_key._tuple(), out_key._tuple()))
    _require(kept is not None, "pruning backward: no forward map for these keys")
    g = grad_out_feat.contiguous()
    _require(g.size(0) == kept.numel(), "grad_out_feat.size(0) == p_map_manager->size(out_key)")
    grad_in = torch.zeros((mgr.size(in_key), g.size(1)), dtype=g.dtype, device=g.device)
    if kept.numel() > 0:
        _lib.check(lib.mk_rows_scatter(_ptr(g), _ptr(kept), kept.numel(), g.size(1) * g.element_size(),
                                       _ptr(grad_in), _stream()))
    return grad_in


# ---------------------------------------------------------------------------
# local pooling (pybind/extern.hpp:215-330; MinkowskiPooling.py)
# ---------------------------------------------------------------------------
def _local_pool_fwd(in_feat, kernel_size, kernel_stride, kernel_dilation, region_type, offset, pooling_mode,
                    in_key, out_key, mgr, is_transpose, expand):
    lib = mgr._lib
    _require(in_feat.is_cuda and in_feat.dim() == 2, "in_feat must be a 2-D CUDA tensor")
    _require(in_feat.size(0) == mgr.size(in_key), "in_feat.size(0) == p_map_manager->size(in_key)")
    in_feat = in_feat.contiguous()
    _resolve_out_key(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                     is_transpose, expand)
    km = _kernel_map(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                     is_transpose, True)
    mode = _POOL_KIND[PoolingMode(int(pooling_mode))]
    c = in_feat.size(1)
    dev = in_feat.device
    out = torch.empty((km.n_out, c), dtype=in_feat.dtype, device=dev)
    if mode == _lib.MK_POOL_MAX:
        aux = torch.empty((km.n_out, c), dtype=torch.int32, device=dev)
        nnz_p, arg_p = None, _ptr(aux)
    else:
        aux = torch.empty(km.n_out, dtype=in_feat.dtype, device=dev)
        nnz_p, arg_p = _ptr(aux), None
    _lib.check(lib.mk_pool_fwd(_dtype_code(in_feat), mode, _ptr(in_feat), _ptr(km.nbr), km.ld, km.n_out, km.K, c,
                               _ptr(out), nnz_p, arg_p, _stream()))
    return out, aux


def _local_pool_bwd(in_feat, grad_out, aux, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                    pooling_mode, in_key, out_key, mgr, is_transpose):
    lib = mgr._lib
    km = _kernel_map(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                     is_transpose, True)
    g = grad_out.contiguous()
    _require(g.size(0) == km.n_out, "grad_out_feat.size(0) == p_map_manager->size(out_key)")
    mode = _POOL_KIND[PoolingMode(int(pooling_mode))]
    inv, ld_in, kflip = _inverse_table(mgr, km)
    if kflip:   # order of k is irrelevant for a pooling reduction
        kflip = 0
    c = g.size(1)
    dx = torch.empty((km.n_in, c), dtype=g.dtype, device=g.device)
    nnz_p = _ptr(aux) if mode != _lib.MK_POOL_MAX else None
    arg_p = _ptr(aux) if mode == _lib.MK_POOL_MAX else None
    _lib.check(lib.mk_pool_bwd(_dtype_code(g), mode, _ptr(g), nnz_p, _ptr(inv), ld_in, km.n_in, km.K, c, arg_p,
                               km.n_out, _ptr(dx), _stream()))
    return dx


def LocalPoolingForwardGPU(in_feat, kernel_size, kernel_stride, kernel_dilation, region_type, offset, pooling_mode,
                           in_key, out_key, mgr):
    """-> (out_feat, num_nonzero); for max pooling the second tensor holds the
    arg-max input rows (pybind/extern.hpp:246-258)."""
    return _local_pool_fwd(in_feat, kernel_size, kernel_stride, kernel_dilation, region_type, offset, pooling_mode,
                           in_key, out_key, mgr, False, False)


def LocalPoolingBackwardGPU(in_feat, grad_out_feat, num_nonzero, kernel_size, kernel_stride, kernel_dilation,
                            region_type, offset, pooling_mode, in_key, out_key, mgr):
    return _local_pool_bwd(in_feat, grad_out_feat, num_nonzero, kernel_size, kernel_stride, kernel_dilation,
                           region_type, offset, pooling_mode, in_key, out_key, mgr, False)


def LocalPoolingTransposeForwardGPU(in_feat, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                                    generate_new_coordinates, pooling_mode, in_key, out_key, mgr):
    return _local_pool_fwd(in_feat, kernel_size, kernel_stride, kernel_dilation, region_type, offset, pooling_mode,
                           in_key, out_key, mgr, True, bool(generate_new_coordinates))


def LocalPoolingTransposeBackwardGPU(in_feat, grad_out_feat, num_nonzero, kernel_size, kernel_stride,
                                     kernel_dilation, region_type, offset, pooling_mode, in_key, out_key, mgr):
    return _local_pool_bwd(in_feat, grad_out_feat, num_nonzero, kernel_size, kernel_stride, kernel_dilation,
                           region_type, offset, pooling_mode, in_key, out_key, mgr, True)


# ---------------------------------------------------------------------------
# global pooling / broadcast (pybind/extern.hpp:332-392)
# ---------------------------------------------------------------------------
def _seg_ws(lib, x, n_batch):
    return torch.empty(int(lib.mk_segment_workspace_bytes(_dtype_code(x), x.size(1), n_batch)),
                       dtype=torch.uint8, device=x.device)


def GlobalPoolingForwardGPU(in_feat, pooling_mode, in_key, out_key, mgr):
    """-> (out [B, C], num_nonzero [B]); out_key becomes the origin map key."""
    lib = mgr._lib
    _require(in_feat.is_cuda and in_feat.dim() == 2, "in_feat must be a 2-D CUDA tensor")
    _require(in_feat.size(0) == mgr.size(in_key), "in_feat.size(0) == p_map_manager->size(in_key)")
    x = in_feat.contiguous()
    if not out_key.is_key_set():
        out_key.set_key(*mgr.origin().get_key())
    rows, nb = mgr._batch_of_row(in_key)
    mode = _POOL_KIND[PoolingMode(int(pooling_mode))]
    c = x.size(1)
    out = torch.empty((nb, c), dtype=x.dtype, device=x.device)
    if mode == _lib.MK_POOL_MAX:
        aux = torch.empty((nb, c), dtype=torch.int32, device=x.device)
        cnt_p, arg_p = None, _ptr(aux)
    else:
        aux = torch.empty(nb, dtype=x.dtype, device=x.device)
        cnt_p, arg_p = _ptr(aux), None
    ws = _seg_ws(lib, x, nb)
    _lib.check(lib.mk_global_pool_fwd(_dtype_code(x), mode, _ptr(x), _ptr(rows), x.size(0), c, nb, _ptr(out),
                                      cnt_p, arg_p, _ptr(ws), _stream()))
    return out, aux


def GlobalPoolingBackwardGPU(in_feat, grad_out_feat, num_nonzero, pooling_mode, in_key, out_key, mgr):
    lib = mgr._lib
    g = grad_out_feat.contiguous()
    rows, nb = mgr._batch_of_row(in_key)
    _require(g.size(0) == nb, "grad_out_feat.size(0) == number of batches")
    mode = _POOL_KIND[PoolingMode(int(pooling_mode))]
    n, c = mgr.size(in_key), g.size(1)
    dx = torch.empty((n, c), dtype=g.dtype, device=g.device)
    cnt_p = _ptr(num_nonzero) if mode != _lib.MK_POOL_MAX else None
    arg_p = _ptr(num_nonzero) if mode == _lib.MK_POOL_MAX else None
    _lib.check(lib.mk_global_pool_bwd(_dtype_code(g), mode, _ptr(g), cnt_p, _ptr(rows), n, c, nb, arg_p, _ptr(dx),
                                      _stream()))
    return dx


def _bcast_mode(mode) -> int:
    return _lib.MK_BCAST_ADD if BroadcastMode(int(mode)) == BroadcastMode.ELEMENTWISE_ADDITON else _lib.MK_BCAST_MUL


def BroadcastForwardGPU(in_feat, in_feat_glob, broadcast_mode, in_key, glob_key, mgr):
    """out[r] = in[r] (+|*) glob[batch(r)] (pybind/extern.hpp:374-380)."""
    lib = mgr._lib
    x, g = in_feat.contiguous(), in_feat_glob.contiguous()
    _require(x.size(0) == mgr.size(in_key), "in_feat.size(0) == p_map_manager->size(in_key)")
    _require(g.size(0) == mgr.size(glob_key), "in_feat_glob.size(0) == p_map_manager->size(glob_key)")
    _require(x.size(1) == g.size(1) and x.dtype == g.dtype, "feature shapes/dtypes must match")
    rows, nb = mgr._batch_of_row(in_key)
    y = torch.empty_like(x)
    _lib.check(lib.mk_broadcast_fwd(_dtype_code(x), _bcast_mode(broadcast_mode), _ptr(x), _ptr(g), _ptr(rows),
                                    x.size(0), x.size(1), _ptr(y), _stream()))
    return y


def BroadcastBackwardGPU(in_feat, in_feat_glob, grad_out_feat, broadcast_mode, in_key, glob_key, mgr):
    """-> (grad_in, grad_glob) (pybind/extern.hpp:382-392)."""
    lib = mgr._lib
    x, g, dy = in_feat.contiguous(), in_feat_glob.contiguous(), grad_out_feat.contiguous()
    rows, nb = mgr._batch_of_row(in_key)
    dx = torch.empty_like(x)
    dg = torch.empty_like(g)
    ws = _seg_ws(lib, x, nb)
    _lib.check(lib.mk_broadcast_bwd(_dtype_code(x), _bcast_mode(broadcast_mode), _ptr(x), _ptr(g), _ptr(dy),
                                    _ptr(rows), x.size(0), x.size(1), nb, _ptr(dx), _ptr(dg), _ptr(ws), _stream()))
    return dx, dg
