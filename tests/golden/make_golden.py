"""Generate tests/golden/hotpath_small.npz.

The reference's native engine cannot be built or imported in this image
(`src/` absent, prebuilt eggs need CPython 3.9/3.11 + CUDA 11), so
these vectors are NOT reference outputs: they are the numpy oracle's outputs on
fixed seeded inputs, frozen so that the oracle, the C/OpenMP port and the CUDA
path are all pinned to the same bytes ("parity unpinned" against upstream ME,
see DESIGN.md §6).  Before writing, the script re-checks the oracle against the
two independent definitions that need no ME at all: a dense grid must equal
torch.nn.functional.conv3d, and insert_and_map must equal numpy.unique.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402


def main():
    rng = np.random.default_rng(20240)
    c = rng.integers(-6, 7, size=(700, 4)).astype(np.int32)
    c[:, 0] = rng.integers(0, 2, size=700)
    c[rng.integers(0, 700, 150)] = c[rng.integers(0, 700, 150)]          # duplicates
    uc, ui, inv = O.insert_and_map(c)
    # independent check: first-occurrence unique
    _, first = np.unique(c, axis=0, return_index=True)
    assert np.array_equal(np.sort(first), ui)

    offs3 = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    km3 = O.kernel_map(uc, uc, offs3)
    sc = O.stride_coords(uc, [2, 2, 2])
    su, _, _ = O.insert_and_map(sc)
    km_s2 = O.kernel_map(uc, su, O.kernel_offsets([2, 2, 2], [1, 1, 1]))   # k2 s2 encoder map
    gen = O.generative_coords(su, O.kernel_offsets([2, 2, 2], [1, 1, 1]))  # octree split of the parents
    x = rng.standard_normal((len(uc), 4))
    w = rng.standard_normal((27, 4, 8)) / np.sqrt(27 * 4)
    dy = rng.standard_normal((len(uc), 8))
    y = O.conv_forward(x, w, km3, len(uc))
    gi, gw = O.conv_backward(x, dy, w, km3)
    keep = rng.random(len(uc)) < 0.6
    pc, pf = O.prune(uc, x, keep)

    # independent check: dense 5^3 grid == conv3d (x fastest in the kernel index)
    g = np.stack(np.meshgrid(np.arange(5), np.arange(5), np.arange(5), indexing="ij"), -1).reshape(-1, 3)
    gc = np.concatenate([np.zeros((125, 1), np.int64), g], 1).astype(np.int32)
    gx = rng.standard_normal((125, 4))
    gy = O.conv_forward(gx, w, O.kernel_map(gc, gc, offs3), 125)
    vol = torch.from_numpy(gx).T.reshape(1, 4, 5, 5, 5)                     # [1, C, X, Y, Z]
    wt = torch.from_numpy(w).reshape(3, 3, 3, 4, 8).permute(4, 3, 2, 1, 0)  # k = ix + 3 (iy + 3 iz) -> [Co, Ci, X, Y, Z]
    ref = torch.nn.functional.conv3d(vol, wt.contiguous(), padding=1)[0].reshape(8, 125).T.numpy()
    assert np.allclose(gy, ref, atol=1e-12)

    def flat(km):
        n = np.array([len(a) for a, _ in km], np.int64)
        return n, np.concatenate([a for a, _ in km]).astype(np.int32), np.concatenate([b for _, b in km]).astype(np.int32)

    n3, in3, out3 = flat(km3)
    n2, in2, out2 = flat(km_s2)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "hotpath_small.npz"),
                        coords=c, unique_coords=uc, unique_index=ui.astype(np.int64), inverse=inv.astype(np.int64),
                        stride2_coords=su, generative_coords=gen, km3_counts=n3, km3_in=in3, km3_out=out3,
                        km_s2_counts=n2, km_s2_in=in2, km_s2_out=out2, x=x, w=w, dy=dy, y=y, grad_in=gi, grad_w=gw,
                        keep=keep, pruned_coords=pc, pruned_feats=pf)
    print("wrote hotpath_small.npz:", len(c), "rows ->", len(uc), "unique,", int(n3.sum()), "pairs")


if __name__ == "__main__":
    main()
