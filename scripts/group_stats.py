"""Fill statistics of the tile-ordered neighbour table at the granularity of TMA tile::gather4
(4 consecutive rows of a 128-row tile under one offset): how many 4-row groups of the non-empty
(tile, offset) pairs are fully valid / partly valid / empty.  North-star cloud (R = 271, 3^3)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import abi  # noqa: E402
from prof_conv import offsets3, sphere  # noqa: E402

c = sphere(int(sys.argv[1]) if len(sys.argv) > 1 else 271)
n = len(c)
tin = abi.insert(c)
nbr, ld, counts = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offsets3())
perm, nbr_o = abi.tile_order(nbr, n, ld, 27, tin["coords_dev"])
t = nbr_o.view(27, ld)[:, : (n // 128) * 128].cpu().numpy() >= 0          # [K, rows]
ntiles = t.shape[1] // 128
v = t.reshape(27, ntiles, 32, 4)
per_group = v.sum(-1)                                                       # [K, tiles, 32] valid rows per group
pair_present = v.reshape(27, ntiles, -1).any(-1)                            # [K, tiles]
g = per_group[pair_present]                                                 # groups of the present pairs
full, empty = (g == 4).sum(), (g == 0).sum()
part = g.size - full - empty
print(f"rows {n}  tiles {ntiles}  present pairs {pair_present.sum()}  fill {t.sum() / (pair_present.sum() * 128):.3f}")
print(f"4-row groups of present pairs: full {full / g.size:.3f}  partial {part / g.size:.3f}  empty {empty / g.size:.3f}")
print(f"rows in full groups / valid rows: {4 * full / t.sum():.3f}")
for w in (8, 16, 32):
    vv = t.reshape(27, ntiles, 128 // w, w)[pair_present]
    s = vv.sum(-1)
    print(f"{w}-row groups: full {(s == w).mean():.3f} partial {((s > 0) & (s < w)).mean():.3f} empty {(s == 0).mean():.3f}")
