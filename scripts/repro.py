"""Small-shape smoke of the tensor-core conv kernels (debug aid): python scripts/repro.py CIN COUT [R]"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import abi
from scripts.prof_conv import sphere, offsets3
cin, cout = int(sys.argv[1]), int(sys.argv[2])
R = int(sys.argv[3]) if len(sys.argv) > 3 else 18
c = sphere(R); n = len(c)
tin = abi.insert(c)
nbr, ld, counts = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offsets3())
g = torch.Generator().manual_seed(0)
x = torch.randn(n, cin, generator=g).bfloat16().cuda()
dy = torch.randn(n, cout, generator=g).bfloat16().cuda()
w = (torch.randn(27, cin, cout, generator=g) / 30).bfloat16().cuda()
for name, fn in [("fwd", lambda: abi.conv_fwd(x, w, nbr, ld, n)),
                 ("dgrad", lambda: abi.conv_fwd(dy, w, nbr, ld, n, transpose_w=1, kflip=1)),
                 ("dgrad_inv", lambda: (lambda iv: abi.conv_fwd(dy, w, iv[0], iv[1], n, transpose_w=1, kflip=0))(abi.kmap_invert(nbr, n, ld, 27, n))),
                 ("wgrad", lambda: abi.conv_wgrad(x, dy, nbr, ld, n, 27))]:
    print(name, "...", flush=True)
    out = fn(); torch.cuda.synchronize()
    print(name, "ok", float(out.float().abs().sum()), "path", abi.last_path(), flush=True)
