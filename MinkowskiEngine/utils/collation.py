"""Batch collation helpers (reference: utils/collation.py:30-263): the batch
index is prepended as column 0 of the coordinates."""
import collections.abc
import logging

import numpy as np
import torch


def _as_tensor(a):
    return torch.from_numpy(a) if isinstance(a, np.ndarray) else a


def batched_coordinates(coords, dtype=torch.int32, device=None):
    """List of [N_i, D] coordinates -> [sum N_i, D+1] with the batch index first."""
    assert isinstance(coords, collections.abc.Sequence), "The coordinates must be a sequence."
    assert all(cs.ndim == 2 for cs in coords), "All coordinates must be in a 2D array."
    dims = {int(cs.shape[1]) for cs in coords}
    assert len(dims) == 1, f"Dimension of the array mismatch. All dimensions: {dims}"
    D = dims.pop()
    if device is None:
        device = coords[0].device if isinstance(coords[0], torch.Tensor) else "cpu"
    assert dtype in [torch.int32, torch.float32], "Only torch.int32, torch.float32 supported for coordinates."
    N = int(sum(len(cs) for cs in coords))
    out = torch.zeros((N, D + 1), dtype=dtype, device=device)
    s = 0
    for b, cs in enumerate(coords):
        cs = _as_tensor(cs)
        if dtype == torch.int32 and cs.dtype.is_floating_point:
            cs = cs.floor()
        n = len(cs)
        out[s:s + n, 1:] = cs.to(dtype)
        out[s:s + n, 0] = b
        s += n
    return out


def sparse_collate(coords, feats, labels=None, dtype=torch.int32, device=None):
    """-> (batched coordinates, concatenated features[, concatenated labels])."""
    assert isinstance(coords, collections.abc.Sequence), "The coordinates must be a sequence of arrays or tensors."
    assert isinstance(feats, collections.abc.Sequence), "The features must be a sequence of arrays or tensors."
    if labels is not None:
        assert isinstance(labels, collections.abc.Sequence), "The labels must be a sequence of arrays or tensors."
    N, Nf = sum(len(c) for c in coords), sum(len(f) for f in feats)
    assert N == Nf, f"Coordinate length {N} != Feature length {Nf}"
    bcoords = batched_coordinates(list(coords), dtype=dtype, device=device)
    feats_batch = torch.cat([_as_tensor(f) for f in feats], 0)
    if labels is None:
        return bcoords, feats_batch
    labels_batch = [_as_tensor(lab) for lab in labels]
    if isinstance(labels_batch[0], torch.Tensor):
        labels_batch = torch.cat(labels_batch, 0)
    return bcoords, feats_batch, labels_batch


def batch_sparse_collate(data, dtype=torch.int32, device=None):
    """collate_fn for a DataLoader yielding (coordinates, features, labels) tuples."""
    return sparse_collate(*list(zip(*data)), dtype=dtype, device=device)


class SparseCollation:
    """collate_fn that truncates the batch once `limit_numpoints` would be exceeded."""

    def __init__(self, limit_numpoints=-1, dtype=torch.int32, device=None):
        self.limit_numpoints = limit_numpoints
        self.dtype = dtype
        self.device = device

    def __call__(self, list_data):
        coords, feats, labels = list(zip(*list_data))
        keep, total = 0, 0
        for c in coords:
            total += c.shape[0]
            if self.limit_numpoints > 0 and total > self.limit_numpoints:
                logging.warning(f"\tCannot fit {sum(len(c) for c in coords)} points into {self.limit_numpoints} "
                                f"points limit. Truncating batch size at {keep} out of {len(coords)}.")
                break
            keep += 1
        return sparse_collate(list(coords[:keep]), list(feats[:keep]), list(labels[:keep]), dtype=self.dtype,
                              device=self.device)
