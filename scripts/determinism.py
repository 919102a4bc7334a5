"""Run the tensor-core kernels several times on the same inputs and count distinct results (debug aid).
    python scripts/determinism.py [R] [sorted|shuffled] [reps]"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import abi
from prof_conv import sphere, offsets3
R = int(sys.argv[1]) if len(sys.argv) > 1 else 120
order = sys.argv[2] if len(sys.argv) > 2 else "shuffled"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 8
c = sphere(R); n = len(c)
if order == "shuffled":
    c = c[np.random.default_rng(3).permutation(n)]
tin = abi.insert(c)
nbr, ld, counts = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offsets3())
g = torch.Generator().manual_seed(0)
x = torch.randn(n, 64, generator=g).bfloat16().cuda()
dy = torch.randn(n, 64, generator=g).bfloat16().cuda()
w = (torch.randn(27, 64, 64, generator=g) / 30).bfloat16().cuda()
mask = abi.tile_mask(nbr, n, ld, 27)
for name, fn in (("fwd", lambda: abi.conv_fwd(x, w, nbr, ld, n, mask=mask)),
                 ("dgrad", lambda: abi.conv_fwd(dy, w, nbr, ld, n, transpose_w=1, kflip=1, mask=mask)),
                 ("wgrad", lambda: abi.conv_wgrad(x, dy, nbr, ld, n, 27, mask=mask))):
    outs = []
    for i in range(reps):
        y = fn(); torch.cuda.synchronize()
        outs.append(y.clone())
    ndiff = sum(int(not torch.equal(outs[0], o)) for o in outs[1:])
    bad = [(int((outs[0] != o).any(dim=-1).reshape(-1).sum()), float((outs[0].float() - o.float()).abs().max())) for o in outs[1:]]
    print(f"{name}: {ndiff}/{reps - 1} runs differ from run 0; (rows differing, max abs diff) = {bad}", flush=True)
