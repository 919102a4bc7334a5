"""Development driver: run the reference's unchanged models on the engine (GPU)
and report errors / step times.  python scripts/run_models.py [vae|unet|minkunet ...] [--res R] [--batch B]"""
import argparse
import os
import sys
import time
import traceback

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from harness import workloads as W


def timed(fn, n=3):
    ts = []
    for _ in range(n):
        torch.cuda.synchronize()
        t = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t)
    return ts


def profile(step, tag):
    from torch.profiler import ProfilerActivity, profile as tprofile
    from mink_octtree_stablediffusion_b200 import backend_ops
    backend_ops.CONV_LOG = []
    with tprofile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    log, backend_ops.CONV_LOG = backend_ops.CONV_LOG, None
    kern = sorted(((e.time_range.start, e.time_range.elapsed_us(), e.name) for e in prof.events()
                   if e.device_type.name == "CUDA" and ("k_conv_tc" in e.name or "k_conv_wgrad_tc" in e.name
                                                         or "k_conv_fwd_simt" in e.name or "k_conv_wgrad_simt" in e.name)))
    print(f"[{tag}] conv launches: {len(kern)} kernels, {len(log)} calls")
    if len(kern) == len(log):
        for (t0, us, name), rec in zip(kern, log):
            kind, n_out, n_in, K, ci, co, dt, ordered = rec
            print(f"  {us:9.1f} us  {kind:5s} n_out={n_out:8d} n_in={n_in:8d} K={K:3d} {ci:4d}->{co:4d} {dt[6:]:8s} "
                  f"{'ordered' if ordered else 'plain  '} {name[:28]}")
    ev = prof.key_averages()
    rows = sorted(((e.device_time_total, e.count, e.key) for e in ev if e.device_time_total > 0 and e.device_type.name == "CUDA"),
                  reverse=True)
    tot = sum(r[0] for r in rows)
    print(f"[{tag}] CUDA kernel time total {tot / 1e3:.2f} ms over {sum(r[1] for r in rows)} launches")
    for t, n, k in rows[:40]:
        print(f"  {t / 1e3:9.3f} ms  {n:5d}x  {k[:110]}")
    cpu = sorted(((e.self_cpu_time_total, e.count, e.key) for e in ev if e.device_type.name == "CPU"), reverse=True)
    print(f"[{tag}] top CPU ops (self time)")
    for t, n, k in cpu[:25]:
        print(f"  {t / 1e3:9.3f} ms  {n:5d}x  {k[:110]}")


def run_vae(a):
    cfg = W.vae_config(resolution=a.res, batch_size=a.batch)
    torch.manual_seed(0)
    model, plm, m = W.load_vae(cfg)
    plm = plm.cuda()
    plm.train()
    opt = plm.configure_optimizers()
    amp = torch.autocast("cuda", dtype=torch.bfloat16, enabled=a.dtype == "bf16")
    coords = torch.from_numpy(W.batched(W.g3_batch(a.res, a.batch))).cuda()
    print(f"[vae] res {a.res} batch {a.batch}: {len(coords)} voxels", flush=True)
    batch = {"coords": coords}

    def step():
        opt.zero_grad(set_to_none=True)
        with amp:
            loss = plm.training_step(batch, 0)
        loss.backward()
        opt.step()
        return loss

    loss = step()
    print("[vae] first step loss", float(loss), flush=True)
    import gc
    gc.collect()
    gc.freeze()
    ts = timed(step, a.steps)
    print("[vae] step times (s):", [round(t, 4) for t in ts], "logged", {k: round(float(v), 5) for k, v in plm._logged.items()})
    if a.profile:
        profile(step, "vae")
    if a.cprofile:
        import cProfile
        import pstats
        pr = cProfile.Profile()
        pr.enable()
        step()
        torch.cuda.synchronize()
        pr.disable()
        st = pstats.Stats(pr)
        st.sort_stats("cumulative").print_stats(45)
        st.sort_stats("tottime").print_stats(30)


def run_minkunet(a):
    mu = W.load_minkunet()
    torch.manual_seed(0)
    net = mu.MinkUNet34C(3, 20, D=3).cuda()
    if a.dtype == "bf16":
        net = net.to(torch.bfloat16)
    import MinkowskiEngine as ME
    from oracle import oracle as O
    c = torch.from_numpy(O.sphere_surface(a.radius, 0, 0)).cuda()
    f = torch.rand(len(c), 3, device="cuda").to(next(net.parameters()).dtype)
    labels = torch.randint(0, 20, (len(c),), device="cuda")
    crit = torch.nn.CrossEntropyLoss()
    print(f"[minkunet] {len(c)} voxels", flush=True)

    def step():
        net.zero_grad(set_to_none=True)
        out = net(ME.SparseTensor(f, c))
        loss = crit(out.F.float(), labels)
        loss.backward()
        return loss

    print("[minkunet] loss", float(step()), flush=True)
    print("[minkunet] step times (s):", [round(t, 4) for t in timed(step, a.steps)])
    if a.profile:
        profile(step, "minkunet")


def run_unet(a):
    cfg = W.diffusion_config(resolution=a.res, batch_size=a.batch, unet_channel=tuple(a.unet_channel))
    d = W.load_unet(cfg)
    import MinkowskiEngine as ME
    torch.manual_seed(0)
    unet = d.UNet(cfg.unet_channel, cfg).cuda()
    if a.dtype == "bf16":
        unet = unet.to(torch.bfloat16)
    dt = next(unet.parameters()).dtype
    # latents: stride-8 coordinates of the G3 shapes
    from oracle import oracle as O
    c = W.batched(W.g3_batch(a.res, a.batch))
    lat = torch.from_numpy(O.stride_coords(c, [8, 8, 8])).cuda()
    x = ME.SparseTensor(torch.randn(len(lat), 4, device="cuda").to(dt), lat, tensor_stride=8)
    print(f"[unet] {len(lat)} latent voxels", flush=True)
    t = torch.randint(0, 1000, (a.batch,), device="cuda")
    noise = torch.randn(len(lat), 4, device="cuda").to(dt)

    def step():
        unet.zero_grad(set_to_none=True)
        out = unet(x, t)
        loss = torch.nn.functional.mse_loss(out.F, noise)
        loss.backward()
        return loss

    print("[unet] loss", float(step()), flush=True)
    print("[unet] step times (s):", [round(t, 4) for t in timed(step, a.steps)])
    if a.profile:
        profile(step, "unet")
    if a.cprofile:
        import cProfile
        import pstats
        pr = cProfile.Profile()
        pr.enable()
        step()
        torch.cuda.synchronize()
        pr.disable()
        st = pstats.Stats(pr)
        st.sort_stats("cumulative").print_stats(70)
        st.sort_stats("tottime").print_stats(40)


if __name__ == "__main__":
    pa = argparse.ArgumentParser()
    pa.add_argument("which", nargs="*", default=["vae"])
    pa.add_argument("--res", type=int, default=64)
    pa.add_argument("--batch", type=int, default=2)
    pa.add_argument("--steps", type=int, default=3)
    pa.add_argument("--radius", type=int, default=40)
    pa.add_argument("--dtype", default="fp32")
    pa.add_argument("--profile", action="store_true")
    pa.add_argument("--cprofile", action="store_true")
    pa.add_argument("--unet_channel", type=int, nargs=4, default=(4, 320, 640, 960))
    a = pa.parse_args()
    rc = 0
    for w in a.which:
        try:
            {"vae": run_vae, "unet": run_unet, "minkunet": run_minkunet}[w](a)
        except Exception:
            rc = 1
            print(f"[{w}] FAILED", flush=True)
            traceback.print_exc()
            try:
                torch.cuda.synchronize()
            except Exception as e:
                print("cuda error after failure:", e)
                break
    sys.exit(rc)
