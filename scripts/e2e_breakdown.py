"""Where the end-to-end step of bench.py spends its time (debug aid)."""
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
coords_d = torch.empty_like(coords_h, device=dev); feats_d = torch.empty_like(feats_h, device=dev)
conv = ME.MinkowskiConvolution(64, 64, kernel_size=3, dimension=3).to(dev).to(torch.bfloat16)
dy = torch.randn(n, 64, device=dev).bfloat16()
grad_h = torch.empty_like(conv.kernel, device="cpu").pin_memory()
def ev(): e = torch.cuda.Event(enable_timing=True); e.record(); return e
for it in range(int(os.environ.get("ITERS", "6"))):
    t0 = time.perf_counter(); e = [ev()]
    coords_d.copy_(coords_h, non_blocking=True); feats_d.copy_(feats_h, non_blocking=True); e.append(ev())
    conv.kernel.grad = None
    x = feats_d.detach().requires_grad_(True)
    s = ME.SparseTensor(x, coords_d); e.append(ev())
    o = conv(s); e.append(ev())
    o.F.backward(dy); e.append(ev())
    grad_h.copy_(conv.kernel.grad, non_blocking=True); e.append(ev())
    torch.cuda.synchronize(); t1 = time.perf_counter()
    names = ["h2d", "SparseTensor(insert)", "conv fwd(+kmap)", "backward", "d2h"]
    print(f"iter {it}: wall {1e3*(t1-t0):.2f} ms | " + " | ".join(f"{nm} {e[i].elapsed_time(e[i+1]):.2f}" for i, nm in enumerate(names)), flush=True)
