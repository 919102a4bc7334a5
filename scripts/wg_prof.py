"""Role wait-time counters of k_conv_wgrad_tc (library built with -DMK_TC_PROF).

    MK_NVCC_EXTRA=-DMK_TC_PROF python -m mink_octtree_stablediffusion_b200.build && python scripts/wg_prof.py
"""
import ctypes as C
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
from mink_octtree_stablediffusion_b200 import _lib  # noqa: E402

lib = _lib.load()
c = sphere(271)
n = len(c)
tin = abi.insert(c)
nbr, ld, counts = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offsets3())
CIN = int(os.environ.get("WG_CIN", "64"))
COUT = int(os.environ.get("WG_COUT", "64"))
x = torch.randn(n, CIN).bfloat16().cuda()
dy = torch.randn(n, COUT).bfloat16().cuda()
perm, nbr_o = abi.tile_order(nbr, n, ld, 27, tin["coords_dev"])   # the product layout
mask_o = abi.tile_mask(nbr_o, n, ld, 27)
names = {0: "mma wait b_full", 1: "mma wait a_full", 2: "mma issue", 3: "mma loop total", 4: "team0 in-panel block", 5: "team0 tile sched", 6: "dy loader wait b_empty", 7: "dy loader index + issue", 24: "dy loader total", 25: "idx lane0 wait i_empty", 26: "idx total", 27: "idx: mask + schedule", 28: "idx: waits + copies (all lanes)", 29: "idx: tile_sched only"}
for t in range(4):
    names.update({8 + 4 * t: f"team{t} wait i_full", 9 + 4 * t: f"team{t} wait a_empty", 10 + 4 * t: f"team{t} issue",
                  11 + 4 * t: f"team{t} loop total"})
for kn in sys.argv[1:] or ["0"]:
    os.environ["MK_WG_KNOCK"] = kn
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        abi.conv_wgrad(x, dy, nbr_o, ld, n, 27, mask=mask_o, dy_rows=perm)
        b.record()
        torch.cuda.synchronize()
    buf = np.zeros(148 * 32, np.uint64)
    lib.mk_debug_wg_prof_read.argtypes = [C.c_void_p, C.c_int]
    assert lib.mk_debug_wg_prof_read(buf.ctypes.data_as(C.c_void_p), len(buf)) == 0
    buf = buf.reshape(148, 32).astype(np.float64)
    print(f"== knock={kn}: {a.elapsed_time(b):.3f} ms; mean over CTAs (kcycles)")
    for k in sorted(names):
        v = buf[:, k]
        print(f"  {names[k]:22s} {v.mean() / 1e3:10.1f}  (min {v.min() / 1e3:.1f} max {v.max() / 1e3:.1f})")
    # loop total per pair group (blockIdx.y = the two halves of buf) and its spread over the tile splits
    if (CIN, COUT) != (64, 64):
        continue
    tot = buf[:, 3].reshape(2, -1) / 1e3
    for y in range(2):
        print(f"  pair group {y}: mma loop total mean {tot[y].mean():.1f} min {tot[y].min():.1f} max {tot[y].max():.1f}; "
              f"first 12 splits {np.round(tot[y][:12], 0)}")
