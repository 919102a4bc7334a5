"""Kernel-map cache and the `*GPU` op functions of the `_C` surface
(pybind/extern.hpp:84-114,151-181,215-428,579-646) on the C-ABI library."""
from __future__ import annotations

import os

import ctypes
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from .backend import (BroadcastMode, ConvolutionMode, CoordinateMapKey, CoordinateMapManagerGPU_c10,
                      PoolingMode, RegionType, _CoordMap, _KernelMap, _POOL_KIND, _dtype_code, _host_i32,
                      _np_ptr, _pad_ld, _ptr, _require, _stream, region_offsets)


# ---------------------------------------------------------------------------
# kernel maps
# ---------------------------------------------------------------------------
def _build_table(mgr, out_map: _CoordMap, in_map: _CoordMap, offsets: np.ndarray, symmetric: bool) -> _KernelMap:
    """nbr[k, o] = row of in_map at out_coords[o] + offsets[k] (or -1)."""
    K = len(offsets)
    ld = _pad_ld(out_map.n)
    dev = out_map.coords.device
    nbr = torch.empty((K, ld), dtype=torch.int32, device=dev)
    offs_h = _host_i32(offsets)
    # per-offset pair counts are not needed by the table form (the CSR export counts on its own)
    _lib.check(mgr._lib.mk_kmap_build(_ptr(out_map.coords), out_map.n, out_map.coords.shape[1], _ptr(in_map.table),
                                      in_map.capacity, _np_ptr(offs_h), K, _ptr(nbr), ld, None, _stream()))
    km = _KernelMap(nbr, ld, in_map.n, out_map.n, K, None, symmetric)
    km.out_coords, km.in_coords = out_map.coords, in_map.coords
    return km


def _inverse_table(mgr, km: _KernelMap):
    """(table over input rows -> output row, ld, kflip).  For a same-map,
    stride-1, origin-symmetric kernel it is the table itself with the offset
    index mirrored, so nothing is built."""
    if km.symmetric:
        return km.nbr, km.ld, 1
    if km._inv is None:
        ld_in = _pad_ld(km.n_in)
        inv = torch.empty((km.K, ld_in), dtype=torch.int32, device=km.nbr.device)
        _lib.check(mgr._lib.mk_kmap_invert(_ptr(km.nbr), km.n_out, km.ld, km.K, _ptr(inv), km.n_in, ld_in, _stream()))
        km._inv = (inv, ld_in)
    return km._inv[0], km._inv[1], 0


def _tile_mask(mgr, table: torch.Tensor, n_rows: int, ld: int, K: int) -> torch.Tensor:
    """Per 128-row tile, the K-bit mask of offsets with at least one neighbour
    (mk_kmap_tile_mask); lets the tensor-core kernels skip empty (tile, offset) pairs."""
    lib = mgr._lib
    W = int(lib.mk_kmap_tile_mask_words(K))
    ntiles = (n_rows + 127) // 128
    mask = torch.empty((max(ntiles, 1), W), dtype=torch.int32, device=table.device)
    if n_rows > 0:
        _lib.check(lib.mk_kmap_tile_mask(_ptr(table), n_rows, ld, K, _ptr(mask), _stream()))
    return mask


def _fwd_mask(mgr, km: _KernelMap) -> torch.Tensor:
    if km._mask is None:
        km._mask = _tile_mask(mgr, km.nbr, km.n_out, km.ld, km.K)
    return km._mask


def _inv_mask(mgr, km: _KernelMap) -> torch.Tensor:
    if km.symmetric:
        return _fwd_mask(mgr, km)
    if km._imask is None:
        inv, ld_in, _ = _inverse_table(mgr, km)
        km._imask = _tile_mask(mgr, inv, km.n_in, ld_in, km.K)
    return km._imask


# Tile ordering (mk_kmap_tile_order): an internal column order of the table for the
# tensor-core kernels.  MK_TILE_ORDER=0 switches it off, MK_TILE_ORDER_MIN is the
# smallest table it is built for (below that the sort costs more than it saves),
# MK_TILE_ORDER_BLOCK the log2 of the locality block.
def _tile_order_cfg():
    import os
    return (os.environ.get("MK_TILE_ORDER", "1") != "0", int(os.environ.get("MK_TILE_ORDER_MIN", "65536")),
            int(os.environ.get("MK_TILE_ORDER_BLOCK", "8")))


def _order_table(mgr, table, n_rows, ld, K, coords):
    lib = mgr._lib
    dev = table.device
    order = torch.empty(max(n_rows, 1), dtype=torch.int32, device=dev)
    out = torch.empty((K, ld), dtype=torch.int32, device=dev)
    ws = torch.empty(int(lib.mk_kmap_tile_order_workspace_bytes(n_rows)), dtype=torch.uint8, device=dev)
    _lib.check(lib.mk_kmap_tile_order(_ptr(table), n_rows, ld, K, _ptr(coords), 0 if coords is None else coords.shape[1],
                                      _tile_order_cfg()[2], _ptr(order), _ptr(out), _ptr(ws), _stream()))
    return order, out, _tile_mask(mgr, out, n_rows, ld, K)


def _want_order(mgr, x, n_rows, K, c_in, c_out, have=False) -> bool:
    """`have`: the ordered form exists already (octree split: written in closed form), so the size
    threshold below which building it does not pay is moot."""
    on, n_min, _ = _tile_order_cfg()
    return (on and (have or n_rows >= n_min) and x.dtype in (torch.bfloat16, torch.float16, torch.float32)
            and bool(mgr._lib.mk_conv_rows_eligible(_dtype_code(x), n_rows, K, c_in, c_out)))


def _ordered(mgr, km: _KernelMap):
    """(order, table, mask) of the forward table in tile order."""
    if km._ordered is None:
        km._ordered = _order_table(mgr, km.nbr, km.n_out, km.ld, km.K, km.out_coords)
    return km._ordered


def _inv_ordered(mgr, km: _KernelMap):
    """Tile-ordered inverse table: (order, table, mask, ld, kflip)."""
    if km.symmetric:
        o, t, m = _ordered(mgr, km)
        return o, t, m, km.ld, 1
    if km._iordered is None:
        inv, ld_in, _ = _inverse_table(mgr, km)
        km._iordered = _order_table(mgr, inv, km.n_in, ld_in, km.K, km.in_coords)
    o, t, m = km._iordered
    return o, t, m, km._inv[1], 0


def _pairs(mgr, km: _KernelMap):
    """CSR export (starts[K+1] host list, in_rows, out_rows); pairs of an offset ascend in the out row."""
    if km._pairs is None:
        lib = mgr._lib
        dev = km.nbr.device
        ws = torch.empty(int(lib.mk_kmap_pairs_workspace_bytes(km.n_out, km.K)), dtype=torch.uint8, device=dev)
        starts = torch.empty(km.K + 1, dtype=torch.int64, device=dev)
        _lib.check(lib.mk_kmap_pairs(_ptr(km.nbr), km.n_out, km.ld, km.K, ctypes.c_void_p(starts.data_ptr()),
                                     None, None, _ptr(ws), _stream()))
        starts_h = starts.cpu().tolist()
        P = int(starts_h[-1])
        in_rows = torch.empty(max(P, 1), dtype=torch.int32, device=dev)
        out_rows = torch.empty(max(P, 1), dtype=torch.int32, device=dev)
        if P > 0:
            _lib.check(lib.mk_kmap_pairs(_ptr(km.nbr), km.n_out, km.ld, km.K, ctypes.c_void_p(starts.data_ptr()),
                                         _ptr(in_rows), _ptr(out_rows), _ptr(ws), _stream()))
        km._pairs = (starts_h, in_rows[:P], out_rows[:P])
    return km._pairs


def _kernel_map(mgr: CoordinateMapManagerGPU_c10, in_key: CoordinateMapKey, out_key: CoordinateMapKey,
                kernel_size: Sequence[int], kernel_stride: Sequence[int], kernel_dilation: Sequence[int],
                region_type, offset: Optional[torch.Tensor], is_transpose: bool, is_pool: bool) -> _KernelMap:
    """Cached kernel map between two coordinate maps.

    Non-transposed: out row u pairs with the in row at u + off_k * dilation * ts_in.
    Transposed (in = coarse, out = fine): out row f pairs with the in row at
    f - off_k * dilation * ts_out, i.e. f = c + off_k * ts_out (upstream ME 0.5.4
    convention recorded in SURVEY.md §8c)."""
    kernel_size = [int(k) for k in kernel_size]
    kernel_stride = [int(s) for s in kernel_stride]
    kernel_dilation = [int(d) for d in kernel_dilation]
    cache_key = _kmap_cache_key(in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                                is_transpose, is_pool)
    km = mgr._kmaps.get(cache_key)
    if km is not None:
        return km
    in_map, out_map = mgr._get(in_key), mgr._get(out_key)
    if not is_transpose:
        scale = [d * t for d, t in zip(kernel_dilation, in_map.tensor_stride)]
        offs = region_offsets(kernel_size, scale, region_type, offset)
        symmetric = (in_key._tuple() == out_key._tuple() and all(s == 1 for s in kernel_stride)
                     and np.array_equal(offs, -offs[::-1]))
        km = _build_table(mgr, out_map, in_map, offs, symmetric)
    else:
        scale = [d * t for d, t in zip(kernel_dilation, out_map.tensor_stride)]
        offs = region_offsets(kernel_size, scale, region_type, offset)
        km = _build_table(mgr, out_map, in_map, -offs, False)
    mgr._kmaps[cache_key] = km
    return km


def _kmap_cache_key(in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset, is_transpose,
                    is_pool):
    return (in_key._tuple(), out_key._tuple(), tuple(int(k) for k in kernel_size), tuple(int(s) for s in kernel_stride),
            tuple(int(d) for d in kernel_dilation), int(region_type), bool(is_transpose), bool(is_pool),
            None if offset is None or offset.numel() == 0 else tuple(offset.reshape(-1).tolist()))


def _children_kernel_map(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset):
    """Transposed kernel map of an octree split (mgr._octree_children): child row i*K + k hangs under
    offset k of parent i, so the table, its (dense) inverse and the tile-ordered form are written in
    closed form (mk_kmap_children) and put into the kernel-map cache - no probes, no sort."""
    in_map, out_map = mgr._get(in_key), mgr._get(out_key)
    K = out_map.n // max(in_map.n, 1)
    dev = in_map.coords.device
    ld_out, ld_in = _pad_ld(out_map.n), _pad_ld(in_map.n)
    nbr = torch.empty((K, ld_out), dtype=torch.int32, device=dev)
    inv = torch.empty((K, ld_in), dtype=torch.int32, device=dev)
    order = torch.empty(out_map.n, dtype=torch.int32, device=dev)
    nbr_o = torch.empty((K, ld_out), dtype=torch.int32, device=dev)
    _lib.check(mgr._lib.mk_kmap_children(in_map.n, K, ld_out, ld_in, _ptr(nbr), _ptr(inv), _ptr(order), _ptr(nbr_o),
                                         _stream()))
    km = _KernelMap(nbr, ld_out, in_map.n, out_map.n, K, None, False)
    km.out_coords, km.in_coords = out_map.coords, in_map.coords
    km._inv = (inv, ld_in)
    km._ordered = (order, nbr_o, _tile_mask(mgr, nbr_o, out_map.n, ld_out, K))
    km._imask = _tile_mask(mgr, inv, in_map.n, ld_in, K)
    km._iordered = (None, inv, km._imask)       # every entry of the inverse is valid: nothing to order
    for is_pool in (False, True):
        mgr._kmaps[_kmap_cache_key(in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                                   True, is_pool)] = km
    return km


def _kernel_map_th(self, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type,
                   offset, is_transpose, is_pool) -> Dict[int, torch.Tensor]:
    """`manager.kernel_map` -> dict{kernel index: int32 [2, n]} with [0] = input
    rows, [1] = output rows (MinkowskiCoordinateManager.py:389-392); offsets
    without pairs are omitted."""
    km = _kernel_map(self, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type,
                     offset, is_transpose, is_pool)
    starts, in_rows, out_rows = _pairs(self, km)
    res = {}
    for k in range(km.K):
        a, b = int(starts[k]), int(starts[k + 1])
        if b > a:
            res[k] = torch.stack([in_rows[a:b], out_rows[a:b]])
    return res


CoordinateMapManagerGPU_c10.kernel_map = _kernel_map_th


class CoordinateMapManagerGPU_default(CoordinateMapManagerGPU_c10):
    """`GPUMemoryAllocatorType.CUDA` variant of the reference
    (pybind/minkowski.cu:57-62).  This engine always allocates through
    PyTorch's caching allocator, so the two classes behave identically."""


class CoordinateMapManagerCPU:
    def __init__(self, *args, **kwargs):
        raise RuntimeError("this engine has no CPU backend (CUDA sm_100a only): create the SparseTensor from "
                           "CUDA coordinates or pass coordinate_map_type=CoordinateMapType.CUDA")


# ---------------------------------------------------------------------------
# convolution
# ---------------------------------------------------------------------------
def _check_conv(in_feat, kernel, in_key, mgr, c_in_axis=1):
    _require(in_feat.is_cuda, "in_feat.is_cuda()")
    _require(in_feat.is_contiguous(), "in_feat.is_contiguous()")
    _require(kernel.is_cuda and kernel.device == in_feat.device, "kernel must be on the feature device")
    _require(in_feat.dim() == 2, "in_feat.dim() == 2")
    _require(kernel.dim() == 3, "kernel.dim() == 3")
    _require(in_feat.size(1) == kernel.size(c_in_axis), "in_feat.size(1) == kernel.size(1)")
    _require(in_feat.dtype == kernel.dtype, "in_feat.scalar_type() == kernel.scalar_type()")
    _require(in_feat.size(0) == mgr.size(in_key), "in_feat.size(0) == p_map_manager->size(in_key)")


CONV_LOG = None      # development aid: set to a list to record (kind, n_out, n_in, K, c_in, c_out, dtype) per launch


_WS_BYTES: Dict[tuple, int] = {}


def _workspace(lib, dtype_code, n_out, n_in, K, c_in, c_out, dev):
    # (the fp32 policy knobs change what the library needs: they are part of the key)
    env = os.environ
    key = (dtype_code, n_out, n_in, K, c_in, c_out, env.get("MK_F32_TC"), env.get("MK_F32_TC_CHAIN"),
           env.get("MK_F32_TERMS"), env.get("MK_TC_NOKSPLIT"), env.get("MK_WG_PREGATHER"), env.get("MK_WG_NB"))
    nbytes = _WS_BYTES.get(key)
    if nbytes is None:
        if len(_WS_BYTES) > 4096:
            _WS_BYTES.clear()
        nbytes = _WS_BYTES[key] = int(lib.mk_conv_workspace_bytes2(dtype_code, n_out, n_in, K, c_in, c_out))
    return torch.empty(nbytes, dtype=torch.uint8, device=dev), nbytes


def _aligned16(*tensors) -> bool:
    return not any(t.data_ptr() & 15 for t in tensors)


def _conv_apply(mgr, x, w, transpose_w, kflip, table, mask, ld, n_rows, K, c_in, c_out, order=None):
    """y[o] = sum_k x[table[k, o]] @ W[k] (W[k]^T when transpose_w); with `order` the
    table is tile-ordered and its column s is output row order[s]."""
    lib = mgr._lib
    y = torch.empty((n_rows, c_out), dtype=x.dtype, device=x.device)
    if n_rows == 0:
        return y
    code = _dtype_code(x)
    if CONV_LOG is not None:
        CONV_LOG.append(("dgrad" if transpose_w else "fwd", n_rows, x.size(0), K, c_in, c_out, str(x.dtype), order is not None))
    ws, nbytes = _workspace(lib, code, n_rows, x.size(0), K, c_in, c_out, x.device)
    _lib.check(lib.mk_conv_fwd_ex(code, _ptr(x), _ptr(w), int(transpose_w), int(kflip), _ptr(table), _ptr(mask), ld,
                                  n_rows, x.size(0), K, c_in, c_out, _ptr(y), _ptr(order), _ptr(ws), nbytes, _stream()))
    return y


def _conv_wgrad(mgr, x, dy, km: _KernelMap, c_in, c_out):
    """dW [K, Cin, Cout]; float32 for 16-bit features (the kernel's own fp32 sum: fp32
    parameters then reduce un-rounded gradients over the ranks), else the feature dtype."""
    lib = mgr._lib
    acc32 = x.dtype in (torch.bfloat16, torch.float16)
    dw = torch.empty((km.K, c_in, c_out), dtype=torch.float32 if acc32 else x.dtype, device=x.device)
    code = _dtype_code(x)
    ws, nbytes = _workspace(lib, code, km.n_out, x.size(0), km.K, c_in, c_out, x.device)
    table, mask, order = km.nbr, None, None
    # (the fp32 FFMA weight-gradient kernel compacts the valid rows of a tile itself and reads dy rows
    # in place: the plain table is the faster form for it; the tensor-core kernel wants the ordered one)
    tc = bool(lib.mk_conv_tc_eligible(code, km.n_out, km.K, c_in, c_out))
    if (km.n_out > 0 and x.size(0) > 0 and _aligned16(x, dy) and tc
            and _want_order(mgr, x, km.n_out, km.K, c_in, c_out, km._ordered is not None)):
        # the ordered table pairs column s with dy row order[s]; the kernel reads dy through it
        order, table, mask = _ordered(mgr, km)
    if mask is None:
        mask = _fwd_mask(mgr, km)
    if CONV_LOG is not None:
        CONV_LOG.append(("wgrad", km.n_out, x.size(0), km.K, c_in, c_out, str(x.dtype), order is not None))
    _lib.check(lib.mk_conv_wgrad_ex(code, _ptr(x), _ptr(dy), _ptr(order), _ptr(table), _ptr(mask), km.ld, km.n_out,
                                    x.size(0), km.K, c_in, c_out, _ptr(dw), int(acc32), _ptr(ws), nbytes, _stream()))
    return dw


def _conv_forward(mgr, x, w, km: _KernelMap):
    c_in, c_out = w.size(1), w.size(2)
    if (km.n_out > 0 and x.size(0) > 0 and _aligned16(x)
            and _want_order(mgr, x, km.n_out, km.K, c_in, c_out, km._ordered is not None)):
        order, table, mask = _ordered(mgr, km)
        return _conv_apply(mgr, x, w, 0, 0, table, mask, km.ld, km.n_out, km.K, c_in, c_out, order)
    return _conv_apply(mgr, x, w, 0, 0, km.nbr, _fwd_mask(mgr, km), km.ld, km.n_out, km.K, c_in, c_out)


def _resolve_out_key(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                     is_transpose, expand):
    if out_key.is_key_set():
        return
    in_map = mgr._get(in_key)
    if not is_transpose:
        if expand:
            scale = [d * t for d, t in zip(kernel_dilation, in_map.tensor_stride)]
            offs = region_offsets(kernel_size, scale, region_type, offset)
            # expanding a stride-1 conv mirrors the region: out = in - off (so that in = out + off)
            k = mgr._stride_region(in_key, -offs, in_map.tensor_stride, True)
            _require(all(s == 1 for s in kernel_stride), "expand_coordinates with stride > 1 is not supported")
        else:
            k = mgr.stride(in_key, kernel_stride)
    else:
        _require(all(t % s == 0 for t, s in zip(in_map.tensor_stride, kernel_stride)),
                 "transposed op: tensor stride must be divisible by the kernel stride")
        out_ts = [t // s for t, s in zip(in_map.tensor_stride, kernel_stride)]
        scale = [d * t for d, t in zip(kernel_dilation, out_ts)]
        offs = region_offsets(kernel_size, scale, region_type, offset)
        k = mgr._octree_children(in_key, offs, out_ts) if expand else None
        if k is not None:
            _children_kernel_map(mgr, in_key, k, kernel_size, kernel_stride, kernel_dilation, region_type, offset)
        else:
            k = mgr._stride_region(in_key, offs, out_ts, expand)
    out_key.set_key(*k.get_key())


def ConvolutionForwardGPU(in_feat, kernel, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                          expand_coordinates, convolution_mode, in_key, out_key, mgr):
    """pybind/extern.hpp:84-98.  `out_key` is mutated in place when unset."""
    _check_conv(in_feat, kernel, in_key, mgr)
    _resolve_out_key(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                     False, bool(expand_coordinates))
    km = _kernel_map(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                     False, False)
    _require(kernel.size(0) == km.K, "kernel.size(0) == kernel_volume")
    return _conv_forward(mgr, in_feat, kernel.contiguous(), km)


def _conv_backward(in_feat, grad_out, kernel, km, mgr):
    grad_out = grad_out.contiguous()
    _require(grad_out.size(0) == km.n_out, "grad_out_feat.size(0) == p_map_manager->size(out_key)")
    _require(grad_out.size(1) == kernel.size(2), "grad_out_feat.size(1) == kernel.size(2)")
    _require(grad_out.dtype == in_feat.dtype, "grad_out_feat.scalar_type() == in_feat.scalar_type()")
    kernel = kernel.contiguous()
    # grad_in[i] = sum_k grad_out[inv[k, i]] @ W[k]^T
    if (km.n_in > 0 and km.n_out > 0 and _aligned16(grad_out)
            and _want_order(mgr, grad_out, km.n_in, km.K, kernel.size(2), kernel.size(1))):
        order, inv, imask, ld_in, kflip = _inv_ordered(mgr, km)
        grad_in = _conv_apply(mgr, grad_out, kernel, 1, kflip, inv, imask, ld_in, km.n_in, km.K,
                              kernel.size(2), kernel.size(1), order)
    else:
        inv, ld_in, kflip = _inverse_table(mgr, km)
        grad_in = _conv_apply(mgr, grad_out, kernel, 1, kflip, inv, _inv_mask(mgr, km), ld_in, km.n_in, km.K,
                              kernel.size(2), kernel.size(1))
    grad_w = _conv_wgrad(mgr, in_feat, grad_out, km, kernel.size(1), kernel.size(2))
    return grad_in, grad_w


def ConvolutionBackwardGPU(in_feat, grad_out_feat, kernel, kernel_size, kernel_stride, kernel_dilation,
                           region_type, offset, convolution_mode, in_key, out_key, mgr):
    """pybind/extern.hpp:100-114 -> (grad_in, grad_kernel)."""
    _check_conv(in_feat, kernel, in_key, mgr)
    km = _kernel_map(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                     False, False)
    return _conv_backward(in_feat, grad_out_feat, kernel, km, mgr)


def ConvolutionTransposeForwardGPU(in_feat, kernel, kernel_size, kernel_stride, kernel_dilation, region_type,
                                   offset, expand_coordinates, convolution_mode, in_key, out_key, mgr):
    """pybind/extern.hpp:151-165 (generative when expand_coordinates)."""
    _check_conv(in_feat, kernel, in_key, mgr)
    _resolve_out_key(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                     True, bool(expand_coordinates))
    km = _kernel_map(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                     True, False)
    _require(kernel.size(0) == km.K, "kernel.size(0) == kernel_volume")
    return _conv_forward(mgr, in_feat, kernel.contiguous(), km)


def ConvolutionTransposeBackwardGPU(in_feat, grad_out_feat, kernel, kernel_size, kernel_stride, kernel_dilation,
                                    region_type, offset, convolution_mode, in_key, out_key, mgr):
    """pybind/extern.hpp:167-181."""
    _check_conv(in_feat, kernel, in_key, mgr)
    km = _kernel_map(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                     True, False)
    return _conv_backward(in_feat, grad_out_feat, kernel, km, mgr)
