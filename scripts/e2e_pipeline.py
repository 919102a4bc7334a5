"""Per-step timeline of bench.py's double-buffered end-to-end loop (debug aid)."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import MinkowskiEngine as ME
from bench import make_cloud, R_FULL
dev = torch.device("cuda:0")
c = make_cloud(R_FULL, 0); n = len(c)
coords_h = torch.from_numpy(c).pin_memory()
feats_h = torch.randn(n, 64).bfloat16().pin_memory()
bufs = [(torch.empty_like(coords_h, device=dev), torch.empty_like(feats_h, device=dev)) for _ in range(2)]
conv = ME.MinkowskiConvolution(64, 64, kernel_size=3, dimension=3).to(dev).to(torch.bfloat16)
dy = torch.randn(n, 64, device=dev).bfloat16()
grad_h = torch.empty_like(conv.kernel, device="cpu").pin_memory()
cs = torch.cuda.Stream(dev)
E = lambda: torch.cuda.Event(enable_timing=True)
N = int(os.environ.get("ITERS", "16"))
PIPE = int(os.environ.get("PIPE", "1"))
c0, c1, m0, m1, m2, m3 = ([E() for _ in range(N)] for _ in range(6))
def issue(i):
    with torch.cuda.stream(cs):
        c0[i].record(cs)
        bufs[i & 1][0].copy_(coords_h, non_blocking=True); bufs[i & 1][1].copy_(feats_h, non_blocking=True)
        c1[i].record(cs)
for rep in range(3):
    base = E(); base.record()
    wall = []
    issue(0)
    for i in range(N):
        t0 = time.perf_counter()
        if PIPE and i + 1 < N: issue(i + 1)
        ms = torch.cuda.current_stream()
        ms.wait_event(c1[i]); m0[i].record()
        conv.kernel.grad = None
        x = bufs[i & 1][1].detach().requires_grad_(True)
        s = ME.SparseTensor(x, bufs[i & 1][0]); m1[i].record()
        o = conv(s); m2[i].record()
        o.F.backward(dy)
        grad_h.copy_(conv.kernel.grad, non_blocking=True); m3[i].record()
        t1 = time.perf_counter()
        ms.synchronize()
        t2 = time.perf_counter()
        del s, o, x
        if not PIPE and i + 1 < N: issue(i + 1)
        wall.append((1e3 * (t1 - t0), 1e3 * (t2 - t0), 1e3 * (time.perf_counter() - t0)))
    torch.cuda.synchronize()
    print(f"rep {rep}: total {base.elapsed_time(m3[N-1]):.2f} ms for {N} steps")
    for i in range(N):
        print(f"  step {i:2d}: copy [{base.elapsed_time(c0[i]):7.2f},{base.elapsed_time(c1[i]):7.2f}] compute start {base.elapsed_time(m0[i]):7.2f} "
              f"insert {m0[i].elapsed_time(m1[i]):.2f} fwd {m1[i].elapsed_time(m2[i]):.2f} bwd {m2[i].elapsed_time(m3[i]):.2f} | host issue {wall[i][0]:.2f} sync {wall[i][1]:.2f} del {wall[i][2]:.2f}")
