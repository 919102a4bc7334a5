"""Role wait-time counters of k_conv_tc (library built with -DMK_TC_PROF).

    MK_NVCC_EXTRA=-DMK_TC_PROF python -m mink_octtree_stablediffusion_b200.build && python scripts/tc_prof.py
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
x = torch.randn(n, 64).bfloat16().cuda()
w = (torch.randn(27, 64, 64) / 40).bfloat16().cuda()
perm, nbr_o = abi.tile_order(nbr, n, ld, 27, tin["coords_dev"])   # the product layout
mask_o = abi.tile_mask(nbr_o, n, ld, 27)
names = {0: "epi wait t_full", 1: "epi work", 2: "idx wait i_empty", 3: "mma wait a_full", 4: "mma wait t_empty",
         5: "mma wait b_full", 6: "mma issue", 7: "mma stages", 24: "mma total", 25: "mma: proxy fence", 26: "mma: umma issue",
         27: "mma: commits", 28: "team0 total", 29: "team0 group fetch+sched", 30: "team0 absent tiles"}
for t in range(4):
    names.update({8 + 4 * t: f"team{t} wait a_empty", 9 + 4 * t: f"team{t} wait i_full", 10 + 4 * t: f"team{t} issue",
                  11 + 4 * t: f"team{t} tiles"})
for cfg in sys.argv[1:] or ["default"]:
    if cfg == "default":     # the plan's own choices
        T_ = B_ = I_ = "-"
        for k_ in ("MK_TC_T", "MK_TC_BSTAGES", "MK_TC_ISLOTS"):
            os.environ.pop(k_, None)
    else:
        T_, B_, I_ = cfg.split(",")
        os.environ["MK_TC_T"], os.environ["MK_TC_BSTAGES"], os.environ["MK_TC_ISLOTS"] = T_, B_, I_
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        abi.conv_fwd_rows(x, w, nbr_o, ld, n, perm, mask=mask_o)
        b.record()
        torch.cuda.synchronize()
    buf = np.zeros(148 * 32, np.uint64)
    lib.mk_debug_prof_read.argtypes = [C.c_void_p, C.c_int]
    assert lib.mk_debug_prof_read(buf.ctypes.data_as(C.c_void_p), len(buf)) == 0
    buf = buf.reshape(148, 32).astype(np.float64)
    print(f"== T={T_} B={B_} I={I_}: {a.elapsed_time(b):.3f} ms; mean over CTAs (kcycles; counts raw)")
    for k in sorted(names):
        v = buf[:, k]
        if "stages" in names[k] or "tiles" in names[k]:
            print(f"  {names[k]:22s} {v.mean():10.0f}")
        else:
            print(f"  {names[k]:22s} {v.mean() / 1e3:10.1f}  (min {v.min() / 1e3:.1f} max {v.max() / 1e3:.1f})")
