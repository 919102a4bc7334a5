"""Integer stages of the north-star layer (hash insert, kernel-map build) through
the raw C-ABI with preallocated buffers: CUDA-event medians + HBM fractions.

    python scripts/prof_int.py [--r 271] [--iters 10] [--factor 2,4] [--shuffle 1]
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scripts"))

from abi import HP, P, S  # noqa: E402
from prof_conv import offsets3, sphere  # noqa: E402
from mink_octtree_stablediffusion_b200 import _lib  # noqa: E402


def med(fn, iters):
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts[1:] if len(ts) > 1 else ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--r", type=int, default=271)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--factor", default="0", help="comma list: table capacity = pow2 >= factor*n (0: mk_hash_capacity)")
    ap.add_argument("--shuffle", type=int, default=0)
    ap.add_argument("--hbm", type=float, default=6458.4, help="GB/s peak for the fractions")
    args = ap.parse_args()
    lib = _lib.load()
    c_np = sphere(args.r)
    if args.shuffle:
        c_np = c_np[np.random.default_rng(1).permutation(len(c_np))]
    c = torch.from_numpy(c_np).cuda()
    n, ncols = c.shape
    offs = np.ascontiguousarray(offsets3())
    K = len(offs)
    ld = (n + 127) // 128 * 128
    for f in [float(v) for v in args.factor.split(",")]:
        cap = int(lib.mk_hash_capacity(n))
        if f > 0:
            cap = 1024
            while cap < f * n:
                cap *= 2
        table = torch.empty(cap * 16, dtype=torch.uint8, device="cuda")
        ws = torch.empty(int(lib.mk_insert_workspace_bytes(n)), dtype=torch.uint8, device="cuda")
        uniq = torch.empty(n, dtype=torch.int32, device="cuda")
        inv = torch.empty(n, dtype=torch.int32, device="cuda")
        cout = torch.empty((n, ncols), dtype=torch.int32, device="cuda")
        meta = torch.zeros(2, dtype=torch.int64, device="cuda")
        nbr = torch.empty((K, ld), dtype=torch.int32, device="cuda")
        counts = torch.zeros(K, dtype=torch.int64, device="cuda")

        def ins():
            _lib.check(lib.mk_map_insert(P(c), n, ncols, P(table), cap, P(uniq), P(inv), P(cout), P(ws),
                                         C.c_void_p(meta.data_ptr()), C.c_void_p(meta.data_ptr() + 8), S()))

        def km():
            _lib.check(lib.mk_kmap_build(P(cout), n, ncols, P(table), cap, HP(offs), K, P(nbr), ld,
                                         C.c_void_p(counts.data_ptr()), S()))

        def km0():
            _lib.check(lib.mk_kmap_build(P(cout), n, ncols, P(table), cap, HP(offs), K, P(nbr), ld, None, S()))

        t_ins = med(ins, args.iters)
        assert int(meta[0]) == n
        t_km = med(km, args.iters)
        t_km0 = med(km0, args.iters)
        Pn = int(counts.sum())
        b_ins, b_km = 84 * n, (16 + 32 * K) * n
        print(f"n={n} cap={cap} load={n / cap:.3f} P/n={Pn / n:.2f} | insert {t_ins * 1e3:.0f} us "
              f"({b_ins / t_ins / 1e6 / args.hbm:.3f} of HBM) | kmap build {t_km * 1e3:.0f} us "
              f"({b_km / t_km / 1e6 / args.hbm:.3f} of HBM at 16+32K B/voxel), without counts {t_km0 * 1e3:.0f} us "
              f"({b_km / t_km0 / 1e6 / args.hbm:.3f})", flush=True)


if __name__ == "__main__":
    main()
