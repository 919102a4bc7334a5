"""CPU oracle: a NumPy restatement of the MinkowskiEngine hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under `oracle/` may be imported by the
product (`mink_octtree_stablediffusion_b200/`, `MinkowskiEngine/`,
`MinkowskiEngineBackend/`); only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` use it, as the checker.

PARITY UNPINNED: the reference ships neither the engine sources (`src/` is
absent), nor a loadable binary, nor tests / golden vectors
(§0.3).  The algorithm lives in the third-party dependency NVIDIA
MinkowskiEngine 0.5.4 (`MinkowskiEngine/__init__.py:25`, no pinned commit).
This file restates its published semantics and is anchored on what the tree
does pin: the op signatures (`pybind/extern.hpp`), the docstring properties
listed in SURVEY.md §4, and ME-independent cross-checks (dense `conv3d`,
float64 gradcheck) in `tests/`.

Every function is independent of the CUDA design (sort / searchsorted on
ravelled keys, no hashing, no neighbour table).
"""
from __future__ import annotations

import itertools
from typing import List, Sequence, Tuple

import numpy as np


# ---------------------------------------------------------------------------
# keys
# ---------------------------------------------------------------------------
def ravel_keys(coords: np.ndarray, lo: np.ndarray, span: np.ndarray) -> np.ndarray:
    """Row-major ravel of int coords [N, D+1] into int64 keys, given per-column
    lower bounds `lo` and extents `span` (order-preserving: lexicographic)."""
    c = coords.astype(np.int64) - lo[None, :]
    key = np.zeros(len(c), dtype=np.int64)
    for d in range(c.shape[1]):
        key = key * int(span[d]) + c[:, d]
    return key


def _bounds(*arrays: np.ndarray, pad: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    cat = np.concatenate([a.reshape(-1, a.shape[-1]) for a in arrays if len(a)], axis=0) \
        if any(len(a) for a in arrays) else np.zeros((1, arrays[0].shape[-1]), np.int64)
    lo = cat.min(axis=0).astype(np.int64) - pad
    hi = cat.max(axis=0).astype(np.int64) + pad
    lo[0] = cat[:, 0].min()
    hi[0] = cat[:, 0].max()
    return lo, hi - lo + 1


# ---------------------------------------------------------------------------
# coordinate maps
# ---------------------------------------------------------------------------
def insert_and_map(coords: np.ndarray) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """`CoordinateManager.insert_and_map` (MinkowskiCoordinateManager.py:153-179).

    Returns (unique_coords, unique_index, inverse_map) with
    `coords[unique_index] == unique_coords`, `unique_coords[inverse_map] == coords`
    (docstring :168-176); the first occurrence of a coordinate is kept and
    `unique_index` ascends (upstream ME 0.5.4 convention)."""
    coords = np.asarray(coords)
    if len(coords) == 0:
        z = np.zeros(0, np.int64)
        return coords.copy(), z, z
    lo, span = _bounds(coords)
    keys = ravel_keys(coords, lo, span)
    _, first, inv_sorted = np.unique(keys, return_index=True, return_inverse=True)
    order = np.argsort(first, kind="stable")          # first-occurrence order
    rank = np.empty_like(order)
    rank[order] = np.arange(len(order))
    unique_index = first[order]
    inverse_map = rank[inv_sorted]
    return coords[unique_index], unique_index.astype(np.int64), inverse_map.astype(np.int64)


def stride_coords(coords: np.ndarray, new_tensor_stride: Sequence[int]) -> np.ndarray:
    """`manager.stride` (pybind/extern.hpp:789; MinkowskiCoordinateManager.py:254-268):
    every spatial coordinate is floored to a multiple of the new tensor stride
    (floor division, negative coordinates allowed — docs/terminology.rst:7),
    then deduplicated.  Row order of the result is first-occurrence order."""
    ts = np.asarray(new_tensor_stride, dtype=np.int64)
    out = coords.astype(np.int64).copy()
    out[:, 1:] = np.floor_divide(out[:, 1:], ts[None, :]) * ts[None, :]
    uniq, _, _ = insert_and_map(out.astype(coords.dtype))
    return uniq


def kernel_offsets(kernel_size: Sequence[int], scale: Sequence[int]) -> np.ndarray:
    """HYPER_CUBE kernel-region offsets [K, D] in coordinate units.

    Odd sizes are centred, even sizes start at 0 (MinkowskiConvolution.py:372-387);
    the first spatial axis varies fastest, k = ix + Kx*(iy + Ky*iz) (upstream ME
    0.5.4 kernel_region iterator; SURVEY §8c assumption).  `scale[d]` is
    dilation[d] * tensor_stride[d]."""
    axes = []
    for ks in kernel_size:
        if ks % 2 == 1:
            axes.append(np.arange(-(ks // 2), ks // 2 + 1))
        else:
            axes.append(np.arange(0, ks))
    D = len(kernel_size)
    offs = []
    for idx in itertools.product(*[range(len(a)) for a in reversed(axes)]):
        idx = idx[::-1]                      # idx[0] is the fastest axis
        offs.append([int(axes[d][idx[d]]) * int(scale[d]) for d in range(D)])
    return np.asarray(offs, dtype=np.int64).reshape(-1, D)


def kernel_volume(kernel_size: Sequence[int]) -> int:
    """MinkowskiKernelGenerator.py:38-102 (HYPER_CUBE): product of sizes."""
    return int(np.prod(kernel_size))


def kernel_map(in_coords: np.ndarray, out_coords: np.ndarray,
               offsets: np.ndarray) -> List[Tuple[np.ndarray, np.ndarray]]:
    """`manager.kernel_map` (pybind/extern.hpp:804; MinkowskiCoordinateManager.py
    :377-421): for every output row u and kernel offset k, the input row at
    coordinate u + offsets[k] if it exists.  Returns per offset
    (in_rows, out_rows) sorted by out row (canonical order)."""
    in_coords = np.asarray(in_coords)
    out_coords = np.asarray(out_coords)
    K = len(offsets)
    if len(in_coords) == 0 or len(out_coords) == 0:
        z = np.zeros(0, np.int64)
        return [(z, z) for _ in range(K)]
    pad = int(np.abs(offsets).max()) if K else 0
    lo, span = _bounds(in_coords, out_coords, pad=pad)
    in_keys = ravel_keys(in_coords, lo, span)
    order = np.argsort(in_keys, kind="stable")
    sorted_keys = in_keys[order]
    result = []
    for k in range(K):
        q = out_coords.astype(np.int64).copy()
        q[:, 1:] += offsets[k][None, :]
        qk = ravel_keys(q, lo, span)
        pos = np.searchsorted(sorted_keys, qk)
        pos_c = np.minimum(pos, len(sorted_keys) - 1)
        hit = sorted_keys[pos_c] == qk
        out_rows = np.nonzero(hit)[0]
        in_rows = order[pos_c[hit]]
        result.append((in_rows.astype(np.int64), out_rows.astype(np.int64)))
    return result


def kernel_map_transpose(in_coords: np.ndarray, out_coords: np.ndarray,
                         offsets: np.ndarray) -> List[Tuple[np.ndarray, np.ndarray]]:
    """Transposed kernel map (`is_transpose=True`, pybind/extern.hpp:151-181):
    the forward map of (out -> in) with roles swapped; `offsets` are scaled by
    the *output* (fine) tensor stride.  Pair k: out = in + offsets[k]."""
    swapped = kernel_map(out_coords, in_coords, offsets)
    res = []
    for fine_rows, coarse_rows in swapped:      # (in_rows of swapped, out_rows of swapped)
        order = np.argsort(fine_rows, kind="stable")
        res.append((coarse_rows[order], fine_rows[order]))
    return res


def generative_coords(in_coords: np.ndarray, offsets: np.ndarray) -> np.ndarray:
    """Output coordinates of `expand_coordinates=True` (generative transposed
    convolution, MinkowskiConvolution.py:617-642; kernel_region_insert in
    SURVEY §2.3): the union over input voxels v and offsets o of v + o."""
    c = in_coords.astype(np.int64)
    allc = (c[:, None, :] + np.concatenate(
        [np.zeros((len(offsets), 1), np.int64), offsets], axis=1)[None, :, :]).reshape(-1, c.shape[1])
    uniq, _, _ = insert_and_map(allc.astype(in_coords.dtype))
    return uniq


def origin_coords(coords: np.ndarray) -> np.ndarray:
    """Origin map: one row (b, 0, .., 0) per batch index (pybind/extern.hpp:790)."""
    o = np.zeros_like(coords)
    o[:, 0] = coords[:, 0]
    uniq, _, _ = insert_and_map(o)
    return uniq


def union_coords(list_of_coords: Sequence[np.ndarray]) -> np.ndarray:
    """`manager.union_map` domain (MinkowskiTensor.py:511-546)."""
    uniq, _, _ = insert_and_map(np.concatenate(list_of_coords, axis=0))
    return uniq


# ---------------------------------------------------------------------------
# feature ops
# ---------------------------------------------------------------------------
def conv_forward(feats: np.ndarray, weight: np.ndarray,
                 kmap: List[Tuple[np.ndarray, np.ndarray]], n_out: int) -> np.ndarray:
    """ConvolutionForward (pybind/extern.hpp:84-98): out[o] = sum_k in[i] @ W[k],
    W [K, Cin, Cout] (orientation fixed by MinkowskiConvolution.py:312-316)."""
    out = np.zeros((n_out, weight.shape[2]), dtype=np.result_type(feats.dtype, weight.dtype))
    for k, (ir, orr) in enumerate(kmap):
        if len(ir):
            np.add.at(out, orr, feats[ir] @ weight[k])
    return out


def conv_backward(feats: np.ndarray, grad_out: np.ndarray, weight: np.ndarray,
                  kmap: List[Tuple[np.ndarray, np.ndarray]]) -> Tuple[np.ndarray, np.ndarray]:
    """ConvolutionBackward (pybind/extern.hpp:100-114): (grad_in, grad_kernel)."""
    gi = np.zeros_like(feats, dtype=np.result_type(feats.dtype, weight.dtype))
    gw = np.zeros_like(weight, dtype=np.result_type(feats.dtype, weight.dtype))
    for k, (ir, orr) in enumerate(kmap):
        if len(ir):
            np.add.at(gi, ir, grad_out[orr] @ weight[k].T)
            gw[k] = feats[ir].T @ grad_out[orr]
    return gi, gw


def prune(coords: np.ndarray, feats: np.ndarray, keep: np.ndarray):
    """PruningForward (pybind/extern.hpp:411-419; MinkowskiPruning.py:93-96):
    rows where keep is true, order preserved."""
    keep = np.asarray(keep).astype(bool)
    return coords[keep], feats[keep]


def prune_backward(grad_out: np.ndarray, keep: np.ndarray, n_in: int) -> np.ndarray:
    """PruningBackward (MinkowskiPruning.py:62-72): zeros with kept rows copied."""
    g = np.zeros((n_in, grad_out.shape[1]), grad_out.dtype)
    g[np.asarray(keep).astype(bool)] = grad_out
    return g


def pool_forward(feats: np.ndarray, kmap, n_out: int, mode: str):
    """LocalPoolingForward (pybind/extern.hpp:215-260; avg divides by the
    per-row cardinality, MinkowskiPooling.py:195-230).  Returns (out, num_nonzero)."""
    C = feats.shape[1]
    nnz = np.zeros(n_out, np.int64)
    if mode == "max":
        out = np.full((n_out, C), -np.inf)
        for ir, orr in kmap:
            if len(ir):
                np.maximum.at(out, orr, feats[ir])
                np.add.at(nnz, orr, 1)
        out[nnz == 0] = 0
        return out.astype(feats.dtype), nnz
    out = np.zeros((n_out, C), np.float64)
    for ir, orr in kmap:
        if len(ir):
            np.add.at(out, orr, feats[ir].astype(np.float64))
            np.add.at(nnz, orr, 1)
    if mode == "avg":
        out = out / np.maximum(nnz, 1)[:, None]
    return out.astype(feats.dtype), nnz


def global_pool(feats: np.ndarray, batch: np.ndarray, n_batch: int, mode: str):
    """GlobalPoolingForward (pybind/extern.hpp:332-352)."""
    out = np.zeros((n_batch, feats.shape[1]), np.float64)
    cnt = np.zeros(n_batch, np.int64)
    np.add.at(out, batch, feats.astype(np.float64))
    np.add.at(cnt, batch, 1)
    if mode == "avg":
        out = out / np.maximum(cnt, 1)[:, None]
    return out.astype(feats.dtype), cnt


# ---------------------------------------------------------------------------
# canonical forms (row order of derived maps is an artefact)
# ---------------------------------------------------------------------------
def canonical_coords(coords: np.ndarray) -> np.ndarray:
    """Rows sorted lexicographically by (b, x, y, z)."""
    coords = np.asarray(coords)
    if len(coords) == 0:
        return coords
    order = np.lexsort(tuple(coords[:, d] for d in reversed(range(coords.shape[1]))))
    return coords[order]


def canonical_kernel_map(kmap, in_coords: np.ndarray, out_coords: np.ndarray):
    """Each offset's pairs as a sorted array of [in coordinate | out coordinate]."""
    res = []
    for ir, orr in kmap:
        ir = np.asarray(ir, dtype=np.int64)
        orr = np.asarray(orr, dtype=np.int64)
        pairs = np.concatenate([in_coords[ir], out_coords[orr]], axis=1) if len(ir) else \
            np.zeros((0, in_coords.shape[1] + out_coords.shape[1]), in_coords.dtype)
        res.append(canonical_coords(pairs))
    return res


def row_permutation(src_coords: np.ndarray, dst_coords: np.ndarray) -> np.ndarray:
    """perm with src_coords[perm] == dst_coords (both hold the same set of rows)."""
    lo, span = _bounds(src_coords, dst_coords)
    sk = ravel_keys(src_coords, lo, span)
    dk = ravel_keys(dst_coords, lo, span)
    order = np.argsort(sk, kind="stable")
    pos = np.searchsorted(sk[order], dk)
    perm = order[pos]
    assert np.array_equal(src_coords[perm], dst_coords), "coordinate sets differ"
    return perm


# ---------------------------------------------------------------------------
# synthetic inputs (SURVEY §8d generators)
# ---------------------------------------------------------------------------
def sphere_surface(R: int, seed: int, batch_index: int = 0) -> np.ndarray:
    """G1: voxelised sphere surface, N ~ 13.65 R^2, int32 [N, 4]."""
    rng = np.random.default_rng(seed)
    v = rng.normal(size=(40 * R * R, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    c = np.floor(v * R + R + 1).astype(np.int32)
    c = np.unique(c, axis=0)
    b = np.full((len(c), 1), batch_index, np.int32)
    return np.concatenate([b, c], axis=1)


def morton_order(coords: np.ndarray) -> np.ndarray:
    """Permutation sorting rows by (batch, Morton code of x,y,z) — the bit order
    pinned by `.ipynb_checkpoints/Untitled1-checkpoint.ipynb` cell 0:
    morton(32,16,0) == 139264 (x bit i -> bit 3i+2, y -> 3i+1, z -> 3i)."""
    c = coords[:, 1:].astype(np.int64)
    c = c - c.min(axis=0, keepdims=True)
    code = np.zeros(len(c), np.int64)
    for i in range(21):
        code |= ((c[:, 0] >> i) & 1) << (3 * i + 2)
        code |= ((c[:, 1] >> i) & 1) << (3 * i + 1)
        code |= ((c[:, 2] >> i) & 1) << (3 * i)
    return np.lexsort((code, coords[:, 0]))
