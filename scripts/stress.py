"""Stress: many launches of the tensor-core kernels over the plans the round added (single accumulator set,
lean gather loop, offsets split over grid.z, pre-gathered dy, shared index slots), every result compared
bit for bit with the first one and every launch timed (a pipeline race shows as a differing row or a
watchdog trap, a missed barrier wake-up as a launch far above the median).
    python scripts/stress.py [reps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import abi  # noqa: E402
from prof_conv import offsets3, sphere  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
ok = True
for R, chans in ((150, ((64, 64), (128, 128), (256, 256), (96, 96), (64, 32))), (12, ((320, 320), (960, 960), (64, 64)))):
    c = sphere(R)
    n = len(c)
    tin = abi.insert(c)
    nbr, ld, counts = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offsets3())
    perm, nbr_o = abi.tile_order(nbr, n, ld, 27, tin["coords_dev"])
    mask_o = abi.tile_mask(nbr_o, n, ld, 27)
    for cin, cout in chans:
        g = torch.Generator().manual_seed(cin + cout)
        x = torch.randn(n, cin, generator=g).bfloat16().cuda()
        dy = torch.randn(n, cout, generator=g).bfloat16().cuda()
        w = (torch.randn(27, cin, cout, generator=g) / 40).bfloat16().cuda()
        runs = {"fwd": lambda: abi.conv_fwd_rows(x, w, nbr_o, ld, n, perm, mask=mask_o),
                "dgrad": lambda: abi.conv_fwd_rows(dy, w, nbr_o, ld, n, perm, transpose_w=1, kflip=1, mask=mask_o),
                "wgrad": lambda: abi.conv_wgrad(x, dy, nbr_o, ld, n, 27, mask=mask_o, dy_rows=perm)}
        for name, fn in runs.items():
            ref = fn()
            torch.cuda.synchronize()
            ts, bad = [], 0
            for _ in range(reps):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                y = fn()
                b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
                bad += int(not torch.equal(y, ref))
            ts = np.array(ts)
            late = ts[3:]          # the first launches of a shape may hold a cudaMalloc of its workspace
            flag = "" if bad == 0 and late.max() < 3 * np.median(ts) + 0.05 else f"   <-- LOOK (launch {3 + int(late.argmax())})"
            ok &= bad == 0
            print(f"n={n:7d} {cin:4d}->{cout:4d} {name:5s}: {bad}/{reps} differ; median {np.median(ts):.3f} ms, max {ts.max():.3f} ms (launch {int(ts.argmax())}){flag}",
                  flush=True)
print("stress", "ok" if ok else "FAILED")
sys.exit(0 if ok else 1)
