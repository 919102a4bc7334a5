"""A few launches of the integer / normalisation kernels at the north-star size, through the C-ABI
(the command the ncu captures of those kernels are taken from).  Prints CUDA-event times."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import abi  # noqa: E402
from mink_octtree_stablediffusion_b200 import backend_norm as BN  # noqa: E402
from oracle import oracle as O  # noqa: E402


def timed(fn, iters=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


c = O.sphere_surface(271, 0, 0).astype(np.int32)
n = len(c)
offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
tin = abi.insert2(c, int64_maps=True)
print(f"n={n}")
print("insert2 (+alloc, sync) ms", timed(lambda: abi.insert2(c, int64_maps=True)))
print("kmap build (+alloc) ms", timed(lambda: abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offs)))
for cch, dt in ((64, torch.bfloat16), (32, torch.bfloat16), (64, torch.float32)):
    x = torch.randn(n, cch, device="cuda").to(dt)
    sh = x[:1].float().contiguous()
    t1 = timed(lambda: BN.seg_sums(x, x, sh, sh, None, 1))
    A = torch.ones(1, cch, device="cuda")
    t2 = timed(lambda: BN.seg_affine(x, None, None, A, None, A))
    by = n * cch * x.element_size()
    print(f"seg_sums   C={cch} {dt}: {t1:.4f} ms  {by / t1 / 1e6:.0f} GB/s")
    print(f"seg_affine C={cch} {dt}: {t2:.4f} ms  {2 * by / t2 / 1e6:.0f} GB/s")
