"""Seeded inputs of the BASELINE single-layer configurations (shared by the golden generator and the
tests): SURVEY 8d generators G1 (sphere surface), torch CPU generators for features and weights."""
import hashlib

import numpy as np
import torch

from oracle import oracle as O


def inputs(name):
    """(coords int32 [N,4], x float32 [N,Cin], w float32 [27,Cin,Cout], dy float32 [N,Cout]); the
    north-star tensors are rounded to bf16 (the values the bf16 kernels see)."""
    if name == "cfg1":
        c = np.concatenate([O.sphere_surface(61, 0, 0), O.sphere_surface(61, 1, 1)]).astype(np.int32)
        cin, cout, rnd = 32, 64, False
    elif name == "north_star":
        c = O.sphere_surface(271, 0, 0).astype(np.int32)
        cin, cout, rnd = 64, 64, True
    else:
        raise KeyError(name)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(len(c), cin, generator=g)
    w = torch.randn(27, cin, cout, generator=g) / float(np.sqrt(27 * cin))
    dy = torch.randn(len(c), cout, generator=g)
    if rnd:
        x, w, dy = (t.bfloat16().float() for t in (x, w, dy))
    return np.ascontiguousarray(c), x.numpy(), w.numpy(), dy.numpy()


def sample_rows(n):
    return np.random.default_rng(7).choice(n, size=64, replace=False)


def kmap_digest(pairs) -> str:
    """SHA-256 over the canonical kernel map: per offset, in rows then out rows (int32), pairs sorted by out row."""
    h = hashlib.sha256()
    for ir, orr in pairs:
        ir, orr = np.asarray(ir, np.int64), np.asarray(orr, np.int64)
        order = np.argsort(orr, kind="stable")
        h.update(ir[order].astype(np.int32).tobytes())
        h.update(orr[order].astype(np.int32).tobytes())
    return h.hexdigest()
