import os, sys
import torch
ROOT = "/root/repo" if os.path.isdir("/root/repo/tests") else os.getcwd()
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import abi
from prof_conv import offsets3, sphere
c = sphere(271); n = len(c)
tin = abi.insert(c)
nbr, ld, counts = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offsets3())
perm, nbr_o = abi.tile_order(nbr, n, ld, 27, tin["coords_dev"])
mask_o = abi.tile_mask(nbr_o, n, ld, 27)
mask_p = abi.tile_mask(nbr, n, ld, 27)
def t(f, it=5):
    f(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / it
for C in (64, 128, 256, 512):
    x = torch.randn(n, C).bfloat16().cuda(); dy = torch.randn(n, C).bfloat16().cuda()
    dyp = dy[perm.long()].contiguous()
    t_rows = t(lambda: abi.conv_wgrad(x, dy, nbr_o, ld, n, 27, mask=mask_o, dy_rows=perm))
    t_pre = t(lambda: abi.conv_wgrad(x, dyp, nbr_o, ld, n, 27, mask=mask_o))
    t_gather = t(lambda: dy[perm.long()])
    t_plain = t(lambda: abi.conv_wgrad(x, dy, nbr, ld, n, 27, mask=mask_p))
    a = abi.conv_wgrad(x, dy, nbr_o, ld, n, 27, mask=mask_o, dy_rows=perm); b = abi.conv_wgrad(x, dyp, nbr_o, ld, n, 27, mask=mask_o)
    print(f"C={C}: dy through order {t_rows:.3f} ms | pre-gathered dy {t_pre:.3f} ms (+ torch gather {t_gather:.3f}) | plain table {t_plain:.3f} ms | equal {torch.equal(a, b)}", flush=True)
    del x, dy, dyp
