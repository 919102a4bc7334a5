"""Host-side mirror of the reference binding module `MinkowskiEngineBackend._C`
(pybind/extern.hpp, pybind/minkowski.cu) on top of the C-ABI library.

Same names, argument order and error behaviour as the pybind module the
reference's Python layer imports (`MinkowskiCommon.py:33`,
`MinkowskiCoordinateManager.py:33-40`): enums, `CoordinateMapKey`,
`CoordinateMapManagerGPU_c10` / `GPU_default`, and the `*GPU` op functions that
`get_minkowski_function` resolves by string (`MinkowskiCommon.py:101-120`).
There is deliberately no `*CPU` function and no CPU manager: north_star rules
out a CPU fallback, so asking for one raises.

All device memory (tables, maps, features, scratch) is allocated with
`torch.empty` on the current device — the caching allocator, i.e. the
reference's default `GPUMemoryAllocatorType.PYTORCH` — and every kernel runs on
PyTorch's current CUDA stream.
"""
from __future__ import annotations

import ctypes
import functools
import enum
import random
import string
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib

# ---------------------------------------------------------------------------
# enums (pybind/extern.hpp:669-741)
# ---------------------------------------------------------------------------


class GPUMemoryAllocatorType(enum.IntEnum):
    PYTORCH = 0
    CUDA = 1


class CUDAKernelMapMode(enum.IntEnum):
    MEMORY_EFFICIENT = 0
    SPEED_OPTIMIZED = 1


class MinkowskiAlgorithm(enum.IntEnum):
    DEFAULT = 0
    MEMORY_EFFICIENT = 1
    SPEED_OPTIMIZED = 2


class CoordinateMapType(enum.IntEnum):
    CPU = 0
    CUDA = 1


class RegionType(enum.IntEnum):
    HYPER_CUBE = 0
    HYPER_CROSS = 1
    CUSTOM = 2


class PoolingMode(enum.IntEnum):
    LOCAL_SUM_POOLING = 0
    LOCAL_AVG_POOLING = 1
    LOCAL_MAX_POOLING = 2
    GLOBAL_SUM_POOLING_DEFAULT = 3
    GLOBAL_AVG_POOLING_DEFAULT = 4
    GLOBAL_MAX_POOLING_DEFAULT = 5
    GLOBAL_SUM_POOLING_KERNEL = 6
    GLOBAL_AVG_POOLING_KERNEL = 7
    GLOBAL_MAX_POOLING_KERNEL = 8
    GLOBAL_SUM_POOLING_PYTORCH_INDEX = 9
    GLOBAL_AVG_POOLING_PYTORCH_INDEX = 10
    GLOBAL_MAX_POOLING_PYTORCH_INDEX = 11


class BroadcastMode(enum.IntEnum):
    ELEMENTWISE_ADDITON = 0
    ELEMENTWISE_MULTIPLICATION = 1


class ConvolutionMode(enum.IntEnum):
    DEFAULT = 0
    DIRECT_GEMM = 1
    COPY_GEMM = 2


_POOL_KIND = {
    PoolingMode.LOCAL_SUM_POOLING: _lib.MK_POOL_SUM,
    PoolingMode.LOCAL_AVG_POOLING: _lib.MK_POOL_AVG,
    PoolingMode.LOCAL_MAX_POOLING: _lib.MK_POOL_MAX,
}
for _m in PoolingMode:
    if "GLOBAL_SUM" in _m.name:
        _POOL_KIND[_m] = _lib.MK_POOL_SUM
    elif "GLOBAL_AVG" in _m.name:
        _POOL_KIND[_m] = _lib.MK_POOL_AVG
    elif "GLOBAL_MAX" in _m.name:
        _POOL_KIND[_m] = _lib.MK_POOL_MAX

_DTYPES = {
    torch.float32: _lib.MK_F32,
    torch.float64: _lib.MK_F64,
    torch.bfloat16: _lib.MK_BF16,
    torch.float16: _lib.MK_F16,
}


def is_cuda_available() -> bool:
    """pybind/extern.hpp:808-814: compiled with CUDA support."""
    return True


def cuda_version() -> int:
    return 12090


def cudart_version() -> int:
    return 12090


def get_gpu_memory_info() -> Tuple[int, int]:
    """pybind/extern.hpp:832-838 -> (free, total) bytes."""
    return tuple(torch.cuda.mem_get_info())


# ---------------------------------------------------------------------------
# helpers
# ---------------------------------------------------------------------------
def _ptr(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else None


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream() -> ctypes.c_void_p:
    """PyTorch's current CUDA stream of the current device (every kernel is launched there).  The raw
    query is ~10x cheaper than building a torch.cuda.Stream object, and this runs once per launch."""
    if _raw_stream is not None:
        return ctypes.c_void_p(_raw_stream(torch.cuda.current_device()))
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _host_i32(values) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(values, dtype=np.int32))


def _np_ptr(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def _require(cond: bool, msg: str) -> None:
    # reference: C++ ASSERT(cond, msg) -> exception -> Python RuntimeError
    if not cond:
        raise RuntimeError(msg)


def _dtype_code(t: torch.Tensor) -> int:
    code = _DTYPES.get(t.dtype)
    _require(code is not None, f"unsupported feature dtype {t.dtype}")
    return code


# ---------------------------------------------------------------------------
# CoordinateMapKey (pybind/extern.hpp:744-764)
# ---------------------------------------------------------------------------
class CoordinateMapKey:
    __slots__ = ("_coordinate_size", "_tensor_stride", "_string_id", "_set")

    def __init__(self, a, b: Optional[str] = None):
        if b is None and isinstance(a, (int, np.integer)):
            self._coordinate_size = int(a)
            self._tensor_stride: List[int] = []
            self._string_id = ""
            self._set = False
        else:
            self._tensor_stride = [int(s) for s in a]
            self._string_id = str(b if b is not None else "")
            self._coordinate_size = len(self._tensor_stride) + 1
            self._set = True

    def is_key_set(self) -> bool:
        return self._set

    def get_coordinate_size(self) -> int:
        return self._coordinate_size

    def get_key(self) -> Tuple[List[int], str]:
        _require(self._set, "CoordinateMapKey: key not set")
        return (list(self._tensor_stride), self._string_id)

    def set_key(self, a, b: Optional[str] = None) -> None:
        if b is None and isinstance(a, tuple) and len(a) == 2 and isinstance(a[1], str):
            a, b = a
        ts = [int(s) for s in a]
        _require(len(ts) + 1 == self._coordinate_size or not self._set and self._coordinate_size in (0, len(ts) + 1),
                 "CoordinateMapKey: tensor stride size mismatch")
        self._tensor_stride, self._string_id = ts, str(b or "")
        self._coordinate_size = len(ts) + 1
        self._set = True

    def get_tensor_stride(self) -> List[int]:
        return list(self._tensor_stride)

    def _tuple(self):
        return (tuple(self._tensor_stride), self._string_id)

    def __hash__(self) -> int:
        return hash(self._tuple())

    def __eq__(self, other) -> bool:
        return isinstance(other, CoordinateMapKey) and self._set == other._set and \
            self._tuple() == other._tuple() and self._coordinate_size == other._coordinate_size

    def __repr__(self) -> str:
        if not self._set:
            return f"coordinate map key:unset, coordinate size:{self._coordinate_size}"
        return f"coordinate map key:{self._tensor_stride}:{self._string_id!r}"


# ---------------------------------------------------------------------------
# device-side objects owned by the manager
# ---------------------------------------------------------------------------
class _CoordMap:
    """coordinates [n, D+1] int32 + the hash table over them."""
    __slots__ = ("coords", "table", "capacity", "n", "tensor_stride", "batch", "on_grid")

    def __init__(self, coords, table, capacity, tensor_stride):
        self.coords = coords
        self.table = table
        self.capacity = capacity
        self.n = coords.shape[0]
        self.tensor_stride = list(tensor_stride)
        self.batch = None      # lazily: (batch_of_row int32 [n], n_batch)
        self.on_grid = None    # every coordinate is a multiple of the tensor stride (None: not checked yet)


class _KernelMap:
    """Output-stationary neighbour table + lazily derived forms."""
    __slots__ = ("nbr", "ld", "n_in", "n_out", "K", "counts", "_inv", "_pairs", "symmetric", "_mask", "_imask",
                 "out_coords", "in_coords", "_ordered", "_iordered")

    def __init__(self, nbr, ld, n_in, n_out, K, counts, symmetric):
        self.nbr, self.ld, self.n_in, self.n_out, self.K = nbr, ld, n_in, n_out, K
        self.counts = counts          # device int64 [K]
        self._inv = None
        self._pairs = None
        self._mask = None             # tile masks of nbr / of the inverse table (lazy)
        self._imask = None
        self.symmetric = symmetric    # inverse table == table with offsets mirrored
        self.out_coords = None        # coordinates of the table's columns / of the inverse table's columns
        self.in_coords = None
        self._ordered = None          # tile-ordered forms (order, table, mask), lazy: see backend_ops._ordered
        self._iordered = None


def _pad_ld(n: int) -> int:
    return max(128, (n + 127) // 128 * 128)


def region_offsets(kernel_size: Sequence[int], scale: Sequence[int], region_type,
                   custom: Optional[torch.Tensor]) -> np.ndarray:
    """Kernel-region offsets [K, D] in coordinate units; kernel index k has the
    first spatial axis fastest; odd sizes centred, even sizes from 0
    (MinkowskiConvolution.py:372-387, MinkowskiKernelGenerator.py:38-102)."""
    D = len(kernel_size)
    region_type = RegionType(int(region_type))
    if region_type == RegionType.CUSTOM:
        _require(custom is not None and custom.numel() > 0, "CUSTOM region needs region_offsets")
        offs = custom.detach().cpu().numpy().astype(np.int64).reshape(-1, D)
        return offs * np.asarray(scale, dtype=np.int64)[None, :]
    return _regular_offsets(tuple(int(k) for k in kernel_size), tuple(int(v) for v in scale), int(region_type))


@functools.lru_cache(maxsize=256)
def _regular_offsets(kernel_size: Tuple[int, ...], scale: Tuple[int, ...], region_type: int) -> np.ndarray:
    D = len(kernel_size)
    region_type = RegionType(region_type)
    axes = [np.arange(-(k // 2), k // 2 + 1) if k % 2 == 1 else np.arange(0, k) for k in kernel_size]
    if region_type == RegionType.HYPER_CUBE:
        grids = np.meshgrid(*axes[::-1], indexing="ij")          # slowest axis first
        offs = np.stack([g.reshape(-1) for g in grids[::-1]], axis=1)
    else:  # HYPER_CROSS: centre, then +- along each axis (MinkowskiKernelGenerator.py:60-78)
        _require(all(k % 2 == 1 for k in kernel_size), "HYPER_CROSS needs odd kernel sizes")
        rows = [[0] * D]
        for d in range(D):
            for v in axes[d]:
                if v != 0:
                    r = [0] * D
                    r[d] = int(v)
                    rows.append(r)
        offs = np.asarray(rows, dtype=np.int64)
    offs = offs.astype(np.int64) * np.asarray(scale, dtype=np.int64)[None, :]
    offs.setflags(write=False)       # cached: callers negate / copy, never write in place
    return offs


# ---------------------------------------------------------------------------
# CoordinateMapManager (pybind/extern.hpp:767-806)
# ---------------------------------------------------------------------------
class CoordinateMapManagerGPU_c10:
    """Owner of coordinate maps, hash tables and cached kernel maps.

    Mirrors `CoordinateMapManager<int32_t, c10_allocator>` as bound at
    pybind/minkowski.cu:57-62.  Not thread-safe (as the reference); one
    instance per Python thread / device."""

    def __init__(self, algorithm: MinkowskiAlgorithm = MinkowskiAlgorithm.DEFAULT,
                 num_threads: int = 0):
        self._lib = _lib.load()
        self.algorithm = algorithm
        self._maps: Dict[tuple, _CoordMap] = {}
        self._kmaps: Dict[tuple, _KernelMap] = {}
        self._stride_maps: Dict[tuple, torch.Tensor] = {}
        self._origin_key: Optional[CoordinateMapKey] = None
        self._device: Optional[torch.device] = None
        self._rng = random.Random(0x5eed)
        self._pending_status: List[torch.Tensor] = []

    # -- internals ----------------------------------------------------------
    def _check_device(self, t: torch.Tensor) -> None:
        _require(t.is_cuda, "coordinates/features must be CUDA tensors (no CPU backend in this engine)")
        if self._device is None:
            self._device = t.device
        _require(t.device == self._device, "all tensors of a manager must live on one device")

    def _get(self, key: CoordinateMapKey) -> _CoordMap:
        _require(isinstance(key, CoordinateMapKey) and key.is_key_set(), "coordinate map key not set")
        m = self._maps.get(key._tuple())
        _require(m is not None, f"coordinate map {key} not found")
        return m

    def _fresh_id(self, tensor_stride: Sequence[int], string_id: str) -> str:
        ts = tuple(int(s) for s in tensor_stride)
        if (ts, string_id) not in self._maps:
            return string_id
        while True:
            cand = string_id + "-" + "".join(self._rng.choice(string.ascii_letters) for _ in range(5)) \
                if string_id else "".join(self._rng.choice(string.ascii_letters) for _ in range(5))
            if (ts, cand) not in self._maps:
                return cand

    def _insert(self, coords: torch.Tensor, want_maps: bool, unique: bool = False):
        """Hash-insert + dedup + compact (mk_map_insert2: three launches, maps written as int64).
        `unique`: the rows are distinct by construction (pruned rows, octree children): no dedup
        pass and no device-to-host read of the count.
        Returns (coords_u, table, cap, unique_index, inverse)."""
        lib = self._lib
        n, ncols = coords.shape
        dev = coords.device
        cap = int(lib.mk_hash_capacity(n))
        table = torch.empty(cap * 16, dtype=torch.uint8, device=dev)
        ws = torch.empty(int(lib.mk_insert2_workspace_bytes(n)), dtype=torch.uint8, device=dev)
        coords_out = coords if unique else torch.empty_like(coords)
        uniq = torch.empty(n, dtype=torch.int64, device=dev) if want_maps else None
        inv = torch.empty(n, dtype=torch.int64, device=dev) if want_maps else None
        meta = torch.empty(2, dtype=torch.int64, device=dev)   # [n_unique, status(int32 in low half)]
        flags = _lib.MK_INSERT_INT64_MAPS | (_lib.MK_INSERT_UNIQUE if unique else 0)
        _lib.check(lib.mk_map_insert2(_ptr(coords), n, ncols, _ptr(table), cap, flags, _ptr(uniq), _ptr(inv),
                                      _ptr(coords_out), _ptr(ws), ctypes.c_void_p(meta.data_ptr()),
                                      ctypes.c_void_p(meta.data_ptr() + 8), _stream()))
        if unique:
            # out-of-range coordinates of derived maps cannot occur (the source map was checked); the
            # status word is still looked at when the map is next synchronised on
            self._pending_status.append(meta)
            return coords_out, table, cap, uniq, inv
        host = meta.cpu()                       # the one D2H sync a new map costs
        n_u = int(host[0].item())
        status = int(host[1].item()) & 0xFFFFFFFF
        self._check_status(status)
        self._drain_status()
        return coords_out[:n_u], table, cap, (uniq[:n_u] if want_maps else None), inv

    @staticmethod
    def _check_status(status: int) -> None:
        if status == 2:
            raise RuntimeError("coordinate out of the packable range: batch in [0, 65534], "
                               "spatial coordinates in [-32768, 32767]")
        _require(status == 0, f"coordinate hash insert failed (status {status})")

    def _drain_status(self) -> None:
        """Status words of the sync-free inserts issued since the last synchronising one."""
        if self._pending_status:
            pend, self._pending_status = self._pending_status, []
            for st in torch.stack(pend).cpu()[:, 1].tolist():
                self._check_status(int(st) & 0xFFFFFFFF)

    def _register(self, key_tuple, coords_u, table, cap) -> _CoordMap:
        m = _CoordMap(coords_u, table, cap, key_tuple[0])
        self._maps[key_tuple] = m
        return m

    def _new_map_from_coords(self, coords: torch.Tensor, tensor_stride, string_id: str,
                             want_maps: bool = False, unique: bool = False):
        coords_u, table, cap, uniq, inv = self._insert(coords, want_maps, unique)
        sid = self._fresh_id(tensor_stride, string_id)
        kt = (tuple(int(s) for s in tensor_stride), sid)
        self._register(kt, coords_u, table, cap)
        return CoordinateMapKey(list(kt[0]), sid), uniq, inv

    # -- public API ---------------------------------------------------------
    def insert_and_map(self, coordinates: torch.Tensor, tensor_stride: Sequence[int],
                       string_id: str = ""):
        """-> (key, (unique_index, inverse_map)); first occurrence kept, unique_index
        ascending (MinkowskiCoordinateManager.py:153-179)."""
        _require(coordinates.dim() == 2, "coordinates must be [N, D+1]")
        _require(coordinates.dtype == torch.int32, "coordinates must be int32 (torch.IntTensor)")
        _require(coordinates.shape[1] == len(tensor_stride) + 1,
                 "coordinates.size(1) must equal len(tensor_stride) + 1")
        self._check_device(coordinates)
        coords = coordinates.contiguous()
        key, uniq, inv = self._new_map_from_coords(coords, tensor_stride, string_id, want_maps=True)
        return key, (uniq, inv)

    def stride(self, in_key: CoordinateMapKey, stride: Sequence[int], string_id: str = "") -> CoordinateMapKey:
        """Strided child map: coordinates floored to multiples of ts*stride and
        de-duplicated; reused when it already exists for this source
        (pybind/extern.hpp:789; MinkowskiCoordinateManager.py:254-268)."""
        m = self._get(in_key)
        stride = [int(s) for s in stride]
        _require(len(stride) == len(m.tensor_stride), "stride size mismatch")
        _require(all(s > 0 for s in stride), "stride must be positive")
        if all(s == 1 for s in stride):
            return in_key
        out_ts = [t * s for t, s in zip(m.tensor_stride, stride)]
        sid = string_id if string_id else in_key.get_key()[1]
        cache_key = (in_key._tuple(), tuple(stride))
        cached = self._stride_maps.get(cache_key)
        if cached is not None:
            return cached
        kt = (tuple(out_ts), sid)
        if kt in self._maps:
            out_key = CoordinateMapKey(out_ts, sid)
        else:
            tmp = torch.empty_like(m.coords)
            ts_h = _host_i32(out_ts)
            _lib.check(self._lib.mk_coords_stride(_ptr(m.coords), m.n, m.coords.shape[1], _np_ptr(ts_h),
                                                  _ptr(tmp), _stream()))
            out_key, _, _ = self._new_map_from_coords(tmp, out_ts, sid)
            self._get(out_key).on_grid = True
        self._stride_maps[cache_key] = out_key
        return out_key

    def _is_on_grid(self, m: _CoordMap) -> bool:
        """All coordinates of the map are multiples of its tensor stride (true by construction for
        strided / generated / pruned maps; checked once, with one device read, for inserted ones)."""
        if m.on_grid is None:
            ts = torch.tensor(m.tensor_stride, dtype=torch.int32, device=m.coords.device).clamp(min=1)
            m.on_grid = bool((m.coords[:, 1:] % ts == 0).all().item()) if m.n else True
        return m.on_grid

    def _octree_children(self, in_key: CoordinateMapKey, offsets: np.ndarray, out_ts: Sequence[int]):
        """Generative expansion whose offsets stay inside the parent cell (kernel == stride: the
        octree split of modules/vae_block.py:279-296): the K children of distinct parents cannot
        collide, so they are inserted without the dedup pass and without reading a count back, and
        child row i*K + k is known to hang under offset k of parent i.  Returns the new key, or
        None when the expansion is not of that form (general path: expand + dedup insert)."""
        m = self._get(in_key)
        offs = np.asarray(offsets, dtype=np.int64)
        in_ts = np.asarray(m.tensor_stride, dtype=np.int64)
        K = len(offs)
        if K < 2 or m.n == 0 or (offs < 0).any() or (offs >= in_ts[None, :]).any() or len(np.unique(offs, axis=0)) != K:
            return None
        if not self._is_on_grid(m):
            return None
        ncols = m.coords.shape[1]
        tmp = torch.empty((m.n * K, ncols), dtype=torch.int32, device=m.coords.device)
        offs_h = _host_i32(offs)
        _lib.check(self._lib.mk_coords_expand(_ptr(m.coords), m.n, ncols, _np_ptr(offs_h), K, _ptr(tmp), _stream()))
        sid = in_key.get_key()[1]
        key, _, _ = self._new_map_from_coords(tmp, out_ts, self._fresh_id(out_ts, sid + "g" if sid else "g"), unique=True)
        self._get(key).on_grid = True
        return key

    def _stride_region(self, in_key: CoordinateMapKey, offsets: np.ndarray, out_ts: Sequence[int],
                       expand: bool) -> CoordinateMapKey:
        """Output map of a transposed op: the cached finer map when it exists
        (non-generative), else the union of in + offsets (generative,
        expand_coordinates=True: MinkowskiConvolution.py:617-642)."""
        m = self._get(in_key)
        sid = in_key.get_key()[1]
        if not expand:
            kt = (tuple(out_ts), sid)
            if kt in self._maps:
                return CoordinateMapKey(list(out_ts), sid)
        K = len(offsets)
        ncols = m.coords.shape[1]
        tmp = torch.empty((m.n * K, ncols), dtype=torch.int32, device=m.coords.device)
        offs_h = _host_i32(offsets)
        _lib.check(self._lib.mk_coords_expand(_ptr(m.coords), m.n, ncols, _np_ptr(offs_h), K, _ptr(tmp), _stream()))
        key, _, _ = self._new_map_from_coords(tmp, out_ts, sid if not expand else self._fresh_id(out_ts, sid + "g" if sid else "g"))
        return key

    def origin(self) -> CoordinateMapKey:
        """Origin map: one row (b, 0, ..., 0) per batch index, tensor stride 0
        (pybind/extern.hpp:790)."""
        if self._origin_key is not None:
            return self._origin_key
        _require(len(self._maps) > 0, "no coordinate map to derive the origin map from")
        first = next(iter(self._maps.values()))
        tmp = torch.empty_like(first.coords)
        _lib.check(self._lib.mk_coords_origin(_ptr(first.coords), first.n, first.coords.shape[1], _ptr(tmp), _stream()))
        D = first.coords.shape[1] - 1
        # rows of the origin map are ordered by batch index (so that row b == batch b)
        coords_u, _, _, _, _ = self._insert(tmp, False)
        order = torch.argsort(coords_u[:, 0].long())
        coords_sorted = coords_u[order].contiguous()
        coords_u2, table, cap, _, _ = self._insert(coords_sorted, False)
        kt = (tuple([0] * D), self._fresh_id([0] * D, ""))
        self._register(kt, coords_u2, table, cap)
        self._origin_key = CoordinateMapKey(list(kt[0]), kt[1])
        return self._origin_key

    def origin_map_size(self) -> int:
        return self._get(self.origin()).n

    def size(self, key: CoordinateMapKey) -> int:
        return self._get(key).n

    def get_coordinates(self, key: CoordinateMapKey) -> torch.Tensor:
        return self._get(key).coords

    def get_coordinate_map_keys(self, tensor_stride: Sequence[int]) -> List[CoordinateMapKey]:
        ts = tuple(int(s) for s in tensor_stride)
        return [CoordinateMapKey(list(k[0]), k[1]) for k in self._maps if k[0] == ts]

    def get_random_string_id(self, tensor_stride: Sequence[int], string_id: str = "") -> CoordinateMapKey:
        ts = [int(s) for s in tensor_stride]
        return CoordinateMapKey(ts, self._fresh_id(ts, string_id or "r"))

    def _batch_of_row(self, key: CoordinateMapKey):
        m = self._get(key)
        if m.batch is None:
            # dense batch ids = row index in the origin map
            om = self._get(self.origin())
            q = torch.zeros_like(m.coords)
            q[:, 0] = m.coords[:, 0]
            rows = torch.empty(m.n, dtype=torch.int32, device=m.coords.device)
            _lib.check(self._lib.mk_map_lookup(_ptr(om.table), om.capacity, _ptr(q), m.n, q.shape[1], _ptr(rows), _stream()))
            m.batch = (rows, om.n)
        return m.batch

    def origin_map(self, key: CoordinateMapKey) -> List[torch.Tensor]:
        """Per batch index, the rows of `key` (pybind/extern.hpp:800)."""
        rows, nb = self._batch_of_row(key)
        r = rows.long()
        # reference returns (origin key, per-batch row lists): MinkowskiTensor.py `_batchwise_row_indices`
        return self.origin(), [torch.nonzero(r == b, as_tuple=False).view(-1) for b in range(nb)]

    def stride_map(self, in_key: CoordinateMapKey, stride_key: CoordinateMapKey) -> torch.Tensor:
        """[2, n] (in row, strided-map row) (pybind/extern.hpp:803)."""
        mi, mo = self._get(in_key), self._get(stride_key)
        tmp = torch.empty_like(mi.coords)
        ts_h = _host_i32(mo.tensor_stride)
        _lib.check(self._lib.mk_coords_stride(_ptr(mi.coords), mi.n, mi.coords.shape[1], _np_ptr(ts_h), _ptr(tmp), _stream()))
        rows = torch.empty(mi.n, dtype=torch.int32, device=mi.coords.device)
        _lib.check(self._lib.mk_map_lookup(_ptr(mo.table), mo.capacity, _ptr(tmp), mi.n, tmp.shape[1], _ptr(rows), _stream()))
        src = torch.arange(mi.n, dtype=torch.int64, device=rows.device)
        ok = rows >= 0
        return torch.stack([src[ok], rows[ok].long()])

    def union_map(self, in_keys: Sequence[CoordinateMapKey], out_key: CoordinateMapKey) -> List[torch.Tensor]:
        """Create the union of the input maps under `out_key` (mutated in place)
        and return, per input, [2, n_i] (input row, union row)
        (pybind/extern.hpp:802; MinkowskiTensor.py:511-546)."""
        maps = [self._get(k) for k in in_keys]
        _require(len(maps) > 0, "union_map needs at least one input")
        ts = maps[0].tensor_stride
        _require(all(m.tensor_stride == ts for m in maps), "union_map: tensor strides differ")
        cat = torch.cat([m.coords for m in maps], dim=0).contiguous()
        key, _, inv = self._new_map_from_coords(cat, ts, "union", want_maps=True)
        out_key.set_key(key.get_key()[0], key.get_key()[1])
        res, start = [], 0
        for m in maps:
            rows = inv[start:start + m.n]
            res.append(torch.stack([torch.arange(m.n, dtype=torch.int64, device=rows.device), rows]))
            start += m.n
        return res

    def __repr__(self) -> str:
        lines = [f"\t{list(k[0])}:{k[1]!r}:\tnrows={m.n}" for k, m in self._maps.items()]
        lines += [f"\tkernel map {k[0][0]}->{k[1][0]} K={v.K}" for k, v in self._kmaps.items()]
        return "\n".join(lines) + "\n"

    def print_coordinate_map(self, key: CoordinateMapKey) -> str:
        m = self._get(key)
        return f"{key}: nrows={m.n} capacity={m.capacity}"
