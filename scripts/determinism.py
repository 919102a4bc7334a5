"""Run the TC conv forward several times on the same inputs and count distinct results (debug aid)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import abi
from scripts.prof_conv import sphere, offsets3
R = 18
c = sphere(R); n = len(c)
tin = abi.insert(c)
nbr, ld, counts = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offsets3())
import ast
for cin, cout in ast.literal_eval(sys.argv[1]) if len(sys.argv) > 1 else [(64, 64), (32, 64), (96, 64), (64, 48), (128, 64), (96, 48), (128, 128), (192, 64)]:
    g = torch.Generator().manual_seed(0)
    x = torch.randn(n, cin, generator=g).bfloat16().cuda()
    w = (torch.randn(27, cin, cout, generator=g) / 30).bfloat16().cuda()
    outs = []
    for i in range(int(os.environ.get('REPS', '8'))):
        y = abi.conv_fwd(x, w, nbr, ld, n); torch.cuda.synchronize()
        outs.append(y.clone())
    ndiff = sum(int(not torch.equal(outs[0], o)) for o in outs[1:])
    maxd = max(float((outs[0].float() - o.float()).abs().max()) for o in outs[1:])
    print(f"cin={cin} cout={cout}: {ndiff}/N runs differ from run 0, max abs diff {maxd:.4g}", flush=True)
