"""North-star layer through the raw C-ABI, a few launches of each kernel — the
command that the ncu captures under profiles/ are taken from.

    python scripts/prof_conv.py [--r 271] [--cin 64] [--cout 64] [--iters 3] [--what fwd,dgrad,wgrad,kmap,insert]

Prints CUDA-event times (median) per kernel so that the same command gives
the un-profiled numbers.  No oracle involved (inputs are synthetic)."""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import abi  # noqa: E402


def sphere(R, seed=0):
    rng = np.random.default_rng(seed)
    v = rng.normal(size=(40 * R * R, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    c = np.unique(np.floor(v * R + R + 1).astype(np.int32), axis=0)
    return np.concatenate([np.zeros((len(c), 1), np.int32), c], 1)


def offsets3():
    r = np.arange(-1, 2)
    # first spatial axis fastest
    return np.array([[x, y, z] for z in r for y in r for x in r], np.int32)


def reorder(c, how):
    """Host-side experiment: the row order the engine's tile ordering would produce."""
    M = 1 << 11
    key = (c[:, 1].astype(np.int64) * M + c[:, 2]) * M + c[:, 3]
    srt = np.argsort(key)
    ks = key[srt]
    pat = np.zeros(len(c), np.uint64)
    for k, (dx, dy, dz) in enumerate(offsets3()):
        q = key + (int(dx) * M + int(dy)) * M + int(dz)
        i = np.clip(np.searchsorted(ks, q), 0, len(c) - 1)
        pat |= (ks[i] == q).astype(np.uint64) << np.uint64(k)
    if how == "pattern":
        return c[np.argsort(pat, kind="stable")]
    b = int(how[1:])
    blk = ((c[:, 1] >> b).astype(np.int64) * M + (c[:, 2] >> b)) * M + (c[:, 3] >> b)
    return c[np.lexsort((pat, blk))]


def _popc(mask):
    return np.unpackbits(mask.cpu().numpy().view(np.uint8)).sum()


def timed(fn, iters):
    """Mean device time of `iters` back-to-back calls after one warm-up (the queue stays
    full, so the host-side cost of a call does not show up as device idle time)."""
    out = fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        out = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--r", type=int, default=271)
    ap.add_argument("--cin", type=int, default=64)
    ap.add_argument("--cout", type=int, default=64)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--what", default="insert,kmap,fwd,dgrad,wgrad")
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--shuffle", type=int, default=0, help="1: random row order instead of sorted")
    ap.add_argument("--dense", type=int, default=0, help="S: dense SxSxS cube instead of the sphere surface")
    ap.add_argument("--order", default="", help="bN: rows sorted by (2^N block, neighbour pattern); morton; pattern (experiment)")
    ap.add_argument("--tile-order", type=int, default=0, help="1: run the conv kernels on the engine's tile-ordered table")
    ap.add_argument("--sweep", default="", help="semicolon list of T,BSTAGES,ISLOTS settings of the conv kernel (dev knobs)")
    args = ap.parse_args()
    what = set(args.what.split(","))
    dt = {"bf16": torch.bfloat16, "f16": torch.float16, "f32": torch.float32}[args.dtype]
    if args.dense:
        S_ = args.dense
        g3 = np.stack(np.meshgrid(np.arange(S_), np.arange(S_), np.arange(S_), indexing="ij"), -1).reshape(-1, 3)
        c = np.concatenate([np.zeros((len(g3), 1), np.int32), g3.astype(np.int32)], 1)
    else:
        c = sphere(args.r)
    if args.shuffle:
        c = c[np.random.default_rng(1).permutation(len(c))]
    if args.order:
        c = reorder(c, args.order)
    n = len(c)
    offs = offsets3()
    t_ins, tin = timed(lambda: abi.insert(c), args.iters if "insert" in what else 1)
    t_km, (nbr, ld, counts) = timed(lambda: abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offs),
                                    args.iters if "kmap" in what else 1)
    P = int(counts.sum())
    print(f"n={n} P={P} P/n={P / n:.2f} insert(+h2d/d2h)={t_ins:.3f} ms kmap(+alloc)={t_km:.3f} ms", flush=True)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(n, args.cin, generator=g).to(dt).cuda()
    dy = torch.randn(n, args.cout, generator=g).to(dt).cuda()
    w = (torch.randn(27, args.cin, args.cout, generator=g) / np.sqrt(27 * args.cin)).to(dt).cuda()
    mask = abi.tile_mask(nbr, n, ld, 27)
    fl = 2.0 * P * args.cin * args.cout
    if args.tile_order:
        t_to, (perm, nbr_o) = timed(lambda: abi.tile_order(nbr, n, ld, 27, tin["coords_dev"]), args.iters)
        mask_o = abi.tile_mask(nbr_o, n, ld, 27)
        print(f"tile order (+alloc) {t_to:.3f} ms; non-empty (tile, offset) pairs {int(_popc(mask))} -> {int(_popc(mask_o))}", flush=True)
        t, _ = timed(lambda: abi.conv_fwd_rows(x, w, nbr_o, ld, n, perm, mask=mask_o), args.iters)
        print(f"fwd   {t:.3f} ms  {fl / t / 1e9:.1f} TFLOP/s useful  (tile-ordered)", flush=True)
        t, _ = timed(lambda: abi.conv_fwd_rows(dy, w, nbr_o, ld, n, perm, transpose_w=1, kflip=1, mask=mask_o), args.iters)
        print(f"dgrad {t:.3f} ms  {fl / t / 1e9:.1f} TFLOP/s useful  (tile-ordered)", flush=True)
        t, _ = timed(lambda: abi.conv_wgrad(x, dy, nbr_o, ld, n, 27, mask=mask_o, dy_rows=perm), args.iters)
        print(f"wgrad {t:.3f} ms  {fl / t / 1e9:.1f} TFLOP/s useful  (tile-ordered, dy read through the order)", flush=True)
        for cfg in [c_ for c_ in args.sweep.split(";") if c_]:
            T_, B_, I_ = cfg.split(",")
            os.environ["MK_TC_T"], os.environ["MK_TC_BSTAGES"], os.environ["MK_TC_ISLOTS"] = T_, B_, I_
            t, _ = timed(lambda: abi.conv_fwd_rows(x, w, nbr_o, ld, n, perm, mask=mask_o), args.iters)
            t2, _ = timed(lambda: abi.conv_fwd_rows(dy, w, nbr_o, ld, n, perm, transpose_w=1, kflip=1, mask=mask_o),
                          args.iters)
            print(f"T={T_} B={B_} I={I_}: fwd {t:.3f} ms dgrad {t2:.3f} ms (tile-ordered)", flush=True)
        return
    for cfg in [c_ for c_ in args.sweep.split(";") if c_]:
        T_, B_, I_ = cfg.split(",")
        os.environ["MK_TC_T"], os.environ["MK_TC_BSTAGES"], os.environ["MK_TC_ISLOTS"] = T_, B_, I_
        t, _ = timed(lambda: abi.conv_fwd(x, w, nbr, ld, n, mask=mask), args.iters)
        t2, _ = timed(lambda: abi.conv_fwd(dy, w, nbr, ld, n, transpose_w=1, kflip=1, c_out=args.cin, mask=mask),
                      args.iters)
        print(f"T={T_} B={B_} I={I_}: fwd {t:.3f} ms dgrad {t2:.3f} ms path={abi.last_path()}", flush=True)
    if args.sweep:
        for k_ in ("MK_TC_T", "MK_TC_BSTAGES", "MK_TC_ISLOTS"):
            os.environ.pop(k_, None)
    if "fwd" in what:
        t, _ = timed(lambda: abi.conv_fwd(x, w, nbr, ld, n, mask=mask), args.iters)
        print(f"fwd   {t:.3f} ms  {fl / t / 1e9:.1f} TFLOP/s useful  path={abi.last_path()}", flush=True)
    if "dgrad" in what:
        t, _ = timed(lambda: abi.conv_fwd(dy, w, nbr, ld, n, transpose_w=1, kflip=1, c_out=args.cin, mask=mask),
                     args.iters)
        print(f"dgrad {t:.3f} ms  {fl / t / 1e9:.1f} TFLOP/s useful  path={abi.last_path()}", flush=True)
    if "wgrad" in what:
        t, _ = timed(lambda: abi.conv_wgrad(x, dy, nbr, ld, n, 27, mask=mask), args.iters)
        print(f"wgrad {t:.3f} ms  {fl / t / 1e9:.1f} TFLOP/s useful  path={abi.last_path()}", flush=True)


if __name__ == "__main__":
    try:
        main()
    except Exception as e:
        import ctypes
        from mink_octtree_stablediffusion_b200 import _lib
        print("FAILED:", str(e)[:100], flush=True)
        w4 = (ctypes.c_uint * 4)()
        rc = _lib.load().mk_debug_hang_words(w4)
        print("watchdog rc", rc, "magic %x block (%d,%d) warp %d smem 0x%x parity %d" %
              (w4[0], w4[1] & 0xFFFF, w4[1] >> 16, w4[2], w4[3] >> 1, w4[3] & 1), flush=True)
        sys.exit(1)
