"""Oracle-backed `*CPU` names of the `_C` surface.  TEST INFRASTRUCTURE ONLY.

The engine has no CPU path (north_star: no CPU fallback; `CoordinateMapManagerCPU`
raises, no `*CPU` op exists).  For parity tests of whole networks the tests
install THIS module's classes/functions under the `*CPU` names of
`MinkowskiEngineBackend._C` (`installed()` context manager): the reference's
unchanged model code then runs a second time on CPU tensors, in float64, on top
of `oracle/oracle.py` (sort / searchsorted maps, per-offset gather-matmul-scatter),
and the GPU engine's coordinates, losses and gradients are compared with it.

Signatures follow pybind/extern.hpp (:84-114 conv, :151-181 transpose, :215-392
pooling / broadcast, :411-428 pruning, :767-806 manager).  Nothing outside
`tests/` may import this file.
"""
from __future__ import annotations

import contextlib
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch

from . import oracle as O


def _np(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().numpy()


def _offsets(kernel_size, scale, region_type, custom) -> np.ndarray:
    kernel_size = [int(k) for k in kernel_size]
    if all(k == 1 for k in kernel_size):
        return np.zeros((1, len(kernel_size)), np.int64)
    if int(region_type) == 0:
        return O.kernel_offsets(kernel_size, scale)
    if int(region_type) == 1:      # HYPER_CROSS: centre, then +- along each axis
        D = len(kernel_size)
        rows = [[0] * D]
        for d in range(D):
            for v in range(-(kernel_size[d] // 2), kernel_size[d] // 2 + 1):
                if v:
                    r = [0] * D
                    r[d] = v * int(scale[d])
                    rows.append(r)
        return np.asarray(rows, np.int64)
    offs = _np(custom).astype(np.int64).reshape(-1, len(kernel_size))
    return offs * np.asarray(scale, np.int64)[None, :]


class CoordinateMapManagerCPU:
    def __init__(self, algorithm=None, num_threads: int = 0):
        from MinkowskiEngineBackend._C import CoordinateMapKey
        self._Key = CoordinateMapKey
        self._maps: Dict[tuple, np.ndarray] = {}
        self._ts: Dict[tuple, List[int]] = {}
        self._kmaps: Dict[tuple, list] = {}
        self._stride_cache: Dict[tuple, object] = {}
        self._origin = None
        self._prune: Dict[tuple, np.ndarray] = {}
        self._n = 0

    # -- helpers -------------------------------------------------------------
    def _coords(self, key) -> np.ndarray:
        return self._maps[key._tuple()]

    def _fresh(self, ts, sid):
        ts = tuple(int(s) for s in ts)
        while (ts, sid) in self._maps:
            self._n += 1
            sid = f"{sid}-o{self._n}" if sid else f"o{self._n}"
        return ts, sid

    def _add(self, coords: np.ndarray, ts, sid):
        ts, sid = self._fresh(ts, sid)
        self._maps[(ts, sid)] = np.ascontiguousarray(coords.astype(np.int32))
        return self._Key(list(ts), sid)

    # -- pybind surface ------------------------------------------------------------
    def insert_and_map(self, coordinates, tensor_stride, string_id=""):
        u, ui, inv = O.insert_and_map(_np(coordinates))
        key = self._add(u, tensor_stride, string_id)
        return key, (torch.from_numpy(ui), torch.from_numpy(inv))

    def stride(self, in_key, stride, string_id=""):
        stride = [int(s) for s in stride]
        if all(s == 1 for s in stride):
            return in_key
        ck = (in_key._tuple(), tuple(stride))
        if ck in self._stride_cache:
            return self._stride_cache[ck]
        ts = [t * s for t, s in zip(in_key.get_tensor_stride(), stride)]
        sid = string_id or in_key.get_key()[1]
        if (tuple(ts), sid) in self._maps:
            key = self._Key(ts, sid)
        else:
            key = self._add(O.stride_coords(self._coords(in_key), ts), ts, sid)
        self._stride_cache[ck] = key
        return key

    def origin(self):
        if self._origin is None:
            first = next(iter(self._maps.values()))
            o = O.origin_coords(first)
            o = o[np.argsort(o[:, 0], kind="stable")]
            self._origin = self._add(o, [0] * (first.shape[1] - 1), "")
        return self._origin

    def origin_map_size(self):
        return len(self._coords(self.origin()))

    def size(self, key):
        return len(self._coords(key))

    def get_coordinates(self, key):
        return torch.from_numpy(self._coords(key))

    def get_coordinate_map_keys(self, tensor_stride):
        ts = tuple(int(s) for s in tensor_stride)
        return [self._Key(list(k[0]), k[1]) for k in self._maps if k[0] == ts]

    def get_random_string_id(self, tensor_stride, string_id=""):
        ts, sid = self._fresh(tensor_stride, string_id or "r")
        return self._Key(list(ts), sid)

    def _batch_rows(self, key):
        ob = self._coords(self.origin())[:, 0]
        b = self._coords(key)[:, 0]
        pos = np.searchsorted(ob, b)
        return pos.astype(np.int64), len(ob)

    def origin_map(self, key):
        rows, nb = self._batch_rows(key)
        return self.origin(), [torch.from_numpy(np.nonzero(rows == b)[0]) for b in range(nb)]

    def stride_map(self, in_key, stride_key):
        ts = stride_key.get_tensor_stride()
        ci = self._coords(in_key).astype(np.int64).copy()
        ci[:, 1:] = np.floor_divide(ci[:, 1:], np.asarray(ts)[None, :]) * np.asarray(ts)[None, :]
        km = O.kernel_map(self._coords(stride_key), ci, np.zeros((1, ci.shape[1] - 1), np.int64))
        ir, orr = km[0]          # (rows of the strided map, rows of in)
        return torch.from_numpy(np.stack([orr, ir]))

    def union_map(self, in_keys, out_key):
        cs = [self._coords(k) for k in in_keys]
        u, _, inv = O.insert_and_map(np.concatenate(cs, axis=0))
        key = self._add(u, in_keys[0].get_tensor_stride(), "union")
        out_key.set_key(*key.get_key())
        res, start = [], 0
        for c in cs:
            res.append(torch.from_numpy(np.stack([np.arange(len(c)), inv[start:start + len(c)]])))
            start += len(c)
        return res

    def _kmap(self, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
              is_transpose):
        ck = (in_key._tuple(), out_key._tuple(), tuple(int(k) for k in kernel_size),
              tuple(int(k) for k in kernel_stride), tuple(int(k) for k in kernel_dilation), int(region_type),
              bool(is_transpose))
        if ck not in self._kmaps:
            ci, co = self._coords(in_key), self._coords(out_key)
            if not is_transpose:
                scale = [d * t for d, t in zip(kernel_dilation, in_key.get_tensor_stride())]
                self._kmaps[ck] = O.kernel_map(ci, co, _offsets(kernel_size, scale, region_type, offset))
            else:
                scale = [d * t for d, t in zip(kernel_dilation, out_key.get_tensor_stride())]
                self._kmaps[ck] = O.kernel_map_transpose(ci, co, _offsets(kernel_size, scale, region_type, offset))
        return self._kmaps[ck]

    def kernel_map(self, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                   is_transpose, is_pool):
        km = self._kmap(in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                        is_transpose)
        return {k: torch.from_numpy(np.stack([ir, orr]).astype(np.int32)) for k, (ir, orr) in enumerate(km) if len(ir)}

    def __repr__(self):
        return "\n".join(f"\t{list(k[0])}:{k[1]!r}:\tnrows={len(v)}" for k, v in self._maps.items()) + "\n"


def _resolve_out(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset, is_transpose,
                 expand):
    if out_key.is_key_set():
        return
    ts_in = in_key.get_tensor_stride()
    sid = in_key.get_key()[1]
    if not is_transpose:
        if expand:
            scale = [d * t for d, t in zip(kernel_dilation, ts_in)]
            offs = _offsets(kernel_size, scale, region_type, offset)
            k = mgr._add(O.generative_coords(mgr._coords(in_key), -offs), ts_in, sid + "g")
        else:
            k = mgr.stride(in_key, kernel_stride)
    else:
        ts_out = [t // s for t, s in zip(ts_in, kernel_stride)]
        if not expand and (tuple(ts_out), sid) in mgr._maps:
            k = mgr._Key(ts_out, sid)
        else:
            scale = [d * t for d, t in zip(kernel_dilation, ts_out)]
            offs = _offsets(kernel_size, scale, region_type, offset)
            k = mgr._add(O.generative_coords(mgr._coords(in_key), offs), ts_out, sid + "g" if expand else sid)
    out_key.set_key(*k.get_key())


def _conv_fwd(x, w, km, n_out):
    out = torch.zeros((n_out, w.shape[2]), dtype=x.dtype)
    for k, (ir, orr) in enumerate(km):
        if len(ir):
            out.index_add_(0, torch.from_numpy(orr), x[torch.from_numpy(ir)] @ w[k])
    return out


def _conv_bwd(x, g, w, km):
    gi = torch.zeros_like(x)
    gw = torch.zeros_like(w)
    for k, (ir, orr) in enumerate(km):
        if len(ir):
            ti, to = torch.from_numpy(ir), torch.from_numpy(orr)
            gi.index_add_(0, ti, g[to] @ w[k].t())
            gw[k] = x[ti].t() @ g[to]
    return gi, gw


def ConvolutionForwardCPU(in_feat, kernel, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                          expand_coordinates, convolution_mode, in_key, out_key, mgr):
    _resolve_out(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset, False,
                 bool(expand_coordinates))
    km = mgr._kmap(in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset, False)
    return _conv_fwd(in_feat, kernel, km, mgr.size(out_key))


def ConvolutionBackwardCPU(in_feat, grad_out_feat, kernel, kernel_size, kernel_stride, kernel_dilation, region_type,
                           offset, convolution_mode, in_key, out_key, mgr):
    km = mgr._kmap(in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset, False)
    return _conv_bwd(in_feat, grad_out_feat, kernel, km)


def ConvolutionTransposeForwardCPU(in_feat, kernel, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                                   expand_coordinates, convolution_mode, in_key, out_key, mgr):
    _resolve_out(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset, True,
                 bool(expand_coordinates))
    km = mgr._kmap(in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset, True)
    return _conv_fwd(in_feat, kernel, km, mgr.size(out_key))


def ConvolutionTransposeBackwardCPU(in_feat, grad_out_feat, kernel, kernel_size, kernel_stride, kernel_dilation,
                                    region_type, offset, convolution_mode, in_key, out_key, mgr):
    km = mgr._kmap(in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset, True)
    return _conv_bwd(in_feat, grad_out_feat, kernel, km)


def PruningForwardCPU(in_feat, keep, in_key, out_key, mgr):
    keep_np = _np(keep).astype(bool).reshape(-1)
    coords = mgr._coords(in_key)[keep_np]
    if not out_key.is_key_set():
        k = mgr._add(coords, in_key.get_tensor_stride(), in_key.get_key()[1] + "p")
        out_key.set_key(*k.get_key())
    mgr._prune[(in_key._tuple(), out_key._tuple())] = keep_np
    return in_feat[torch.from_numpy(keep_np)]


def PruningBackwardCPU(grad_out_feat, in_key, out_key, mgr):
    keep_np = mgr._prune[(in_key._tuple(), out_key._tuple())]
    g = torch.zeros((len(keep_np), grad_out_feat.shape[1]), dtype=grad_out_feat.dtype)
    g[torch.from_numpy(keep_np)] = grad_out_feat
    return g


def _pool_mode(mode) -> str:
    name = getattr(mode, "name", str(mode))
    return "max" if "MAX" in name else ("avg" if "AVG" in name else "sum")


def _local_pool_fwd(in_feat, kernel_size, kernel_stride, kernel_dilation, region_type, offset, pooling_mode, in_key,
                    out_key, mgr, is_transpose, expand):
    _resolve_out(mgr, in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset, is_transpose,
                 expand)
    km = mgr._kmap(in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset, is_transpose)
    n_out, c = mgr.size(out_key), in_feat.shape[1]
    mode = _pool_mode(pooling_mode)
    if mode == "max":
        out = torch.full((n_out, c), -float("inf"), dtype=in_feat.dtype)
        arg = torch.full((n_out, c), -1, dtype=torch.int32)
        for ir, orr in km:
            if len(ir):
                ti, to = torch.from_numpy(ir), torch.from_numpy(orr)
                v = in_feat[ti]
                better = v > out[to]
                out[to] = torch.where(better, v, out[to])
                arg[to] = torch.where(better, ti.to(torch.int32)[:, None].expand_as(v), arg[to])
        out[arg < 0] = 0
        return out, arg
    out = torch.zeros((n_out, c), dtype=in_feat.dtype)
    nnz = torch.zeros(n_out, dtype=in_feat.dtype)
    for ir, orr in km:
        if len(ir):
            to = torch.from_numpy(orr)
            out.index_add_(0, to, in_feat[torch.from_numpy(ir)])
            nnz.index_add_(0, to, torch.ones(len(ir), dtype=in_feat.dtype))
    if mode == "avg":
        out = out / nnz.clamp(min=1)[:, None]
    return out, nnz


def _local_pool_bwd(in_feat, grad_out, aux, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                    pooling_mode, in_key, out_key, mgr, is_transpose):
    km = mgr._kmap(in_key, out_key, kernel_size, kernel_stride, kernel_dilation, region_type, offset, is_transpose)
    mode = _pool_mode(pooling_mode)
    dx = torch.zeros((mgr.size(in_key), grad_out.shape[1]), dtype=grad_out.dtype)
    if mode == "max":
        rows = aux.long()
        ok = rows >= 0
        cols = torch.arange(grad_out.shape[1])[None, :].expand_as(rows)
        dx.index_put_((rows[ok], cols[ok]), grad_out[ok], accumulate=True)
        return dx
    g = grad_out / aux.clamp(min=1)[:, None] if mode == "avg" else grad_out
    for ir, orr in km:
        if len(ir):
            dx.index_add_(0, torch.from_numpy(ir), g[torch.from_numpy(orr)])
    return dx


def LocalPoolingForwardCPU(in_feat, kernel_size, kernel_stride, kernel_dilation, region_type, offset, pooling_mode,
                           in_key, out_key, mgr):
    return _local_pool_fwd(in_feat, kernel_size, kernel_stride, kernel_dilation, region_type, offset, pooling_mode,
                           in_key, out_key, mgr, False, False)


def LocalPoolingBackwardCPU(in_feat, grad_out_feat, num_nonzero, kernel_size, kernel_stride, kernel_dilation,
                            region_type, offset, pooling_mode, in_key, out_key, mgr):
    return _local_pool_bwd(in_feat, grad_out_feat, num_nonzero, kernel_size, kernel_stride, kernel_dilation,
                           region_type, offset, pooling_mode, in_key, out_key, mgr, False)


def LocalPoolingTransposeForwardCPU(in_feat, kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                                    generate_new_coordinates, pooling_mode, in_key, out_key, mgr):
    return _local_pool_fwd(in_feat, kernel_size, kernel_stride, kernel_dilation, region_type, offset, pooling_mode,
                           in_key, out_key, mgr, True, bool(generate_new_coordinates))


def LocalPoolingTransposeBackwardCPU(in_feat, grad_out_feat, num_nonzero, kernel_size, kernel_stride,
                                     kernel_dilation, region_type, offset, pooling_mode, in_key, out_key, mgr):
    return _local_pool_bwd(in_feat, grad_out_feat, num_nonzero, kernel_size, kernel_stride, kernel_dilation,
                           region_type, offset, pooling_mode, in_key, out_key, mgr, True)


def GlobalPoolingForwardCPU(in_feat, pooling_mode, in_key, out_key, mgr):
    if not out_key.is_key_set():
        out_key.set_key(*mgr.origin().get_key())
    rows, nb = mgr._batch_rows(in_key)
    rows_t = torch.from_numpy(rows)
    mode = _pool_mode(pooling_mode)
    c = in_feat.shape[1]
    if mode == "max":
        out = torch.full((nb, c), -float("inf"), dtype=in_feat.dtype)
        arg = torch.zeros((nb, c), dtype=torch.int32)
        for b in range(nb):
            sel = torch.nonzero(rows_t == b).view(-1)
            if len(sel):
                v, i = in_feat[sel].max(dim=0)
                out[b], arg[b] = v, sel[i].to(torch.int32)
        return out, arg
    out = torch.zeros((nb, c), dtype=in_feat.dtype).index_add_(0, rows_t, in_feat)
    cnt = torch.zeros(nb, dtype=in_feat.dtype).index_add_(0, rows_t, torch.ones(len(rows), dtype=in_feat.dtype))
    if mode == "avg":
        out = out / cnt.clamp(min=1)[:, None]
    return out, cnt


def GlobalPoolingBackwardCPU(in_feat, grad_out_feat, num_nonzero, pooling_mode, in_key, out_key, mgr):
    rows, nb = mgr._batch_rows(in_key)
    rows_t = torch.from_numpy(rows)
    mode = _pool_mode(pooling_mode)
    if mode == "max":
        dx = torch.zeros((len(rows), grad_out_feat.shape[1]), dtype=grad_out_feat.dtype)
        cols = torch.arange(grad_out_feat.shape[1])[None, :].expand(nb, -1)
        dx.index_put_((num_nonzero.long(), cols), grad_out_feat, accumulate=True)
        return dx
    g = grad_out_feat / num_nonzero.clamp(min=1)[:, None] if mode == "avg" else grad_out_feat
    return g[rows_t]


def BroadcastForwardCPU(in_feat, in_feat_glob, broadcast_mode, in_key, glob_key, mgr):
    rows, _ = mgr._batch_rows(in_key)
    g = in_feat_glob[torch.from_numpy(rows)]
    return in_feat + g if int(broadcast_mode) == 0 else in_feat * g


def BroadcastBackwardCPU(in_feat, in_feat_glob, grad_out_feat, broadcast_mode, in_key, glob_key, mgr):
    rows, nb = mgr._batch_rows(in_key)
    rows_t = torch.from_numpy(rows)
    if int(broadcast_mode) == 0:
        dx, dgr = grad_out_feat, grad_out_feat
    else:
        dx, dgr = grad_out_feat * in_feat_glob[rows_t], grad_out_feat * in_feat
    dg = torch.zeros_like(in_feat_glob).index_add_(0, rows_t, dgr)
    return dx.clone(), dg


_NAMES = [n for n in dir() if n.endswith("CPU")]


@contextlib.contextmanager
def installed():
    """Install the `*CPU` names into `MinkowskiEngineBackend._C` (and the alias
    the front-end's manager module holds) for the duration of a test."""
    import MinkowskiEngineBackend._C as _C
    saved = {n: getattr(_C, n, None) for n in _NAMES}
    g = globals()
    for n in _NAMES:
        setattr(_C, n, g[n])
    try:
        yield
    finally:
        for n, v in saved.items():
            if v is None:
                delattr(_C, n)
            else:
                setattr(_C, n, v)
