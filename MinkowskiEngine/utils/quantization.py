"""Voxelisation helpers (reference: utils/quantization.py:31-363).

`sparse_quantize` de-duplicates discretised coordinates keeping the first
occurrence.  On CUDA inputs/devices it runs on the engine's GPU coordinate hash;
for host-side data preparation (numpy / CPU tensors in dataloader workers, the
reference's CPU-manager path) it is plain NumPy — data loading, not the hot path.
"""
from collections.abc import Sequence

import numpy as np
import torch


def fnv_hash_vec(arr):
    """FNV64-1A over the columns of an int array."""
    assert arr.ndim == 2
    arr = arr.copy().astype(np.uint64, copy=False)
    hashed = np.uint64(14695981039346656037) * np.ones(arr.shape[0], dtype=np.uint64)
    for j in range(arr.shape[1]):
        hashed *= np.uint64(1099511628211)
        hashed = np.bitwise_xor(hashed, arr[:, j])
    return hashed


def ravel_hash_vec(arr):
    """Ravel the (shifted-to-zero) coordinates into one uint64 key per row."""
    assert arr.ndim == 2
    arr = arr.copy()
    arr -= arr.min(0)
    arr = arr.astype(np.uint64, copy=False)
    arr_max = arr.max(0).astype(np.uint64) + 1
    keys = np.zeros(arr.shape[0], dtype=np.uint64)
    for j in range(arr.shape[1] - 1):
        keys += arr[:, j]
        keys *= arr_max[j + 1]
    keys += arr[:, -1]
    return keys


def _first_occurrence_maps(discrete: np.ndarray):
    keys = ravel_hash_vec(discrete.astype(np.int64))
    _, first, inv_sorted = np.unique(keys, return_index=True, return_inverse=True)
    order = np.argsort(first, kind="stable")
    rank = np.empty_like(order)
    rank[order] = np.arange(len(order))
    return first[order], rank[inv_sorted.reshape(-1)]


def _auto_floor(array):
    assert isinstance(array, (np.ndarray, torch.Tensor)), "array must be either np.array or torch.Tensor."
    return np.floor(array) if isinstance(array, np.ndarray) else torch.floor(array)


def _is_cuda_device(device) -> bool:
    if isinstance(device, torch.device):
        return device.type == "cuda"
    if isinstance(device, str):
        if device == "cpu":
            return False
        if "cuda" in device:
            return True
    raise ValueError("Invalid device. Only `cpu`, `cuda` or torch.device supported.")


def unique_coordinate_map(coordinates: torch.Tensor, tensor_stride=1):
    """-> (unique_map, inverse_map) of integer coordinates [N, D+1] (batch first)."""
    assert coordinates.ndim == 2, "Coordinates must be a matrix"
    assert isinstance(coordinates, torch.Tensor)
    if coordinates.is_cuda:
        from ..MinkowskiCoordinateManager import CoordinateManager
        mgr = CoordinateManager(D=coordinates.shape[1] - 1)
        _, (umap, imap) = mgr.insert_and_map(coordinates.int(), tensor_stride)
        return umap, imap
    umap, imap = _first_occurrence_maps(coordinates.numpy())
    return torch.from_numpy(umap), torch.from_numpy(imap)


def sparse_quantize(coordinates, features=None, labels=None, ignore_label=-100, return_index=False,
                    return_inverse=False, return_maps_only=False, quantization_size=None, device="cpu"):
    """Voxelise: floor(coordinates / quantization_size) and keep one point per voxel.

    unique_coords[inverse_map] == discrete_coords and
    discrete_coords[unique_map] == unique_coords (reference docstring :190-195)."""
    assert isinstance(coordinates, (np.ndarray, torch.Tensor)), "Coords must be either np.array or torch.Tensor."
    use_label, use_feat = labels is not None, features is not None
    assert coordinates.ndim == 2, \
        "The coordinates must be a 2D matrix. The shape of the input is " + str(coordinates.shape)
    if return_inverse:
        assert return_index, "return_reverse must be set with return_index"
    if use_feat:
        assert features.ndim == 2
        assert coordinates.shape[0] == features.shape[0]
    if use_label:
        assert coordinates.shape[0] == len(labels)
    dimension = coordinates.shape[1]
    if quantization_size is not None:
        if isinstance(quantization_size, (Sequence, np.ndarray, torch.Tensor)):
            assert len(quantization_size) == dimension, "Quantization size and coordinates size mismatch."
            q = np.array([i for i in quantization_size]) if isinstance(coordinates, np.ndarray) \
                else torch.Tensor([i for i in quantization_size])
            discrete = _auto_floor(coordinates / q)
        elif np.isscalar(quantization_size):
            discrete = _auto_floor(coordinates) if quantization_size == 1 else _auto_floor(coordinates / quantization_size)
        else:
            raise ValueError("Not supported type for quantization_size.")
    else:
        discrete = _auto_floor(coordinates)
    discrete = discrete.astype(np.int32) if isinstance(discrete, np.ndarray) else discrete.int()
    is_np = isinstance(discrete, np.ndarray)

    # CUDA coordinates (or device="cuda") are de-duplicated by the engine's GPU coordinate hash
    # (first occurrence kept, ascending: the same maps as the host path)
    on_gpu = _is_cuda_device(device) or (not is_np and discrete.is_cuda)
    if on_gpu:
        import MinkowskiEngineBackend._C as MEB
        dev_coords = (torch.from_numpy(discrete) if is_np else discrete)
        dev_coords = dev_coords if dev_coords.is_cuda else dev_coords.to(device)
        mgr = MEB.CoordinateMapManagerGPU_c10()
        _, (unique_map, inverse_map) = mgr.insert_and_map(dev_coords, [1] * (dimension - 1), "")
        if is_np or not discrete.is_cuda:
            unique_map, inverse_map = unique_map.cpu(), inverse_map.cpu()
        if is_np:
            unique_map, inverse_map = unique_map.numpy(), inverse_map.numpy()
    else:
        umap, imap = _first_occurrence_maps(discrete if is_np else discrete.cpu().numpy())
        if is_np:
            unique_map, inverse_map = umap, imap
        else:
            unique_map = torch.from_numpy(umap).to(discrete.device)
            inverse_map = torch.from_numpy(imap).to(discrete.device)

    if use_label and not is_np and discrete.is_cuda:
        # label voting on the device: a voxel whose points disagree on the label gets ignore_label
        lab = labels.to(discrete.device).reshape(-1)
        first_lab = lab[unique_map]
        conflict = torch.zeros(len(unique_map), dtype=torch.int32, device=lab.device)
        conflict.index_put_((inverse_map,), (lab != first_lab[inverse_map]).int(), accumulate=True)
        colabels = torch.where(conflict > 0, torch.full_like(first_lab, ignore_label), first_lab)
        ret = [discrete[unique_map]]
        if use_feat:
            ret.append(features[unique_map])
        ret.append(colabels)
    elif use_label:
        # a voxel whose points disagree on the label gets ignore_label (reference quantize_label_*)
        lab = labels if is_np else labels.cpu().numpy()
        lab = np.asarray(lab).reshape(-1)
        n_u = len(unique_map)
        inv = inverse_map if is_np else inverse_map.cpu().numpy()
        first_lab = lab[unique_map if is_np else unique_map.cpu().numpy()]
        conflict = np.zeros(n_u, dtype=bool)
        np.logical_or.at(conflict, inv, lab != first_lab[inv])
        colabels = np.where(conflict, ignore_label, first_lab).astype(lab.dtype)
        if not is_np:
            colabels = torch.from_numpy(colabels).to(labels.device)
        ret = [discrete[unique_map]]
        if use_feat:
            ret.append(features[unique_map])
        ret.append(colabels)
    else:
        if return_maps_only:
            return (unique_map, inverse_map) if return_inverse else unique_map
        ret = [discrete[unique_map]]
        if use_feat:
            ret.append(features[unique_map])
    if return_index:
        ret.append(unique_map)
    if return_inverse:
        ret.append(inverse_map)
    return ret[0] if len(ret) == 1 else tuple(ret)
