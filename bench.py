#!/usr/bin/env python
"""bench.py — north-star benchmark: MinkowskiConvolution fwd+bwd voxels/s on a
1 M-voxel sphere-surface cloud, 3x3x3, 64->64 channels, bf16 (BASELINE.json).

  python bench.py --gpus N --steps K --warmup W            (this engine, CUDA)
  python bench.py --impl reference --gpus N --steps K ...  (CPU baseline arm)

A "step" is one pass of the hot path over one batch of synthetic input:
conv forward + data gradient + weight gradient with the kernel map cached (the
reference caches it per (in,out,kernel) triple, docs/guides.md:7-17); the
kernel-map build and hash insert are timed separately and reported beside it
(`stages`).  N>1: one replica of the workload per GPU (weak scaling), the only
collective is the NCCL all-reduce of the weight gradient.

Prints ONE JSON line (rank 0).  See DESIGN.md §Measurement for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "sparse_conv_fwd_bwd_voxels_per_sec"
UNIT = "voxels/s"
R_FULL = 271          # sphere radius -> N ~ 1.00 M voxels
C_IN = C_OUT = 64
KS = 3


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d.get("bf16_tflops_sustained"),
                    source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, source="fallback")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([t.strip() for t in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i] == "Active" for r in self.rows)]
        busy = [s for s in sm if s > 0.6 * (max(mx) if mx else 1)] or sm
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def bind_rank_to_cpus(local: int, world: int):
    """Give every rank its own slice of the host cores that are local to its GPU (nvidia-smi topo -m,
    column 'CPU Affinity'): the e2e leg is bound by the host side (pinned-memory reads, the Python launch
    thread), and eight ranks floating over the same cores take turns.  Returns a description for the JSON."""
    info = {"cpus_total": os.cpu_count()}
    try:
        allowed = sorted(os.sched_getaffinity(0))
        rows = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout.splitlines()
        head = next(r for r in rows if "CPU Affinity" in r)
        cols = [c.strip() for c in head.split("\t")]
        ci = cols.index("CPU Affinity")
        aff = {}
        for r in rows:
            cells = [c.strip() for c in r.split("\t")]
            if cells and cells[0].startswith("GPU") and len(cells) > ci:
                cpus = []
                for part in cells[ci].split(","):
                    a, _, b = part.partition("-")
                    cpus += list(range(int(a), int(b or a) + 1))
                aff[int(cells[0][3:])] = [c for c in cpus if c in allowed]
        mine = aff.get(local) or allowed
        sharing = sorted(g for g, c in aff.items() if c == mine and g < world) or [local]
        per = max(1, len(mine) // len(sharing))
        i = sharing.index(local) if local in sharing else 0
        mine = mine[i * per:(i + 1) * per] or mine
        os.sched_setaffinity(0, mine)
        info.update(gpu_cpu_affinity=cells_to_str(aff.get(local, [])), bound_to=cells_to_str(mine))
    except Exception as e:          # no topology information: leave the affinity alone
        info["bind_error"] = f"{type(e).__name__}: {e}"
    return info


def cells_to_str(cpus):
    return f"{cpus[0]}-{cpus[-1]} ({len(cpus)})" if cpus else ""


def make_cloud(R: int, seed: int, batch_index: int = 0) -> np.ndarray:
    """G1 sphere-surface generator, int32 [N, 4], batch first."""
    rng = np.random.default_rng(seed)
    v = rng.normal(size=(40 * R * R, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    c = np.unique(np.floor(v * R + R + 1).astype(np.int32), axis=0)
    return np.concatenate([np.full((len(c), 1), batch_index, np.int32), c], axis=1)


# ---------------------------------------------------------------------------
# CPU arm: oracle/cpu_ref (C + OpenMP restatement of the ME CPU backend)
# ---------------------------------------------------------------------------
def cpu_sample(R: int, reps: int, seed: int = 0):
    """Time conv fwd+bwd (kernel map cached, like the GPU step) on a sphere of
    radius R, fp32 (the reference CPU backend has no bf16).  -> dict"""
    from oracle import cpu_ref as CR
    from oracle import oracle as O
    c = make_cloud(R, seed)
    n = len(c)
    offs = O.kernel_offsets([KS] * 3, [1, 1, 1])
    t0 = time.perf_counter()
    # every host thread this process may use (torchrun exports OMP_NUM_THREADS=1 for its workers; ME's own
    # default is min(cpu_count, 16-20), __init__.py:34-46)
    CR.set_num_threads(len(os.sched_getaffinity(0)))
    cmap = CR.CoordMap(c)
    t1 = time.perf_counter()
    km = CR.KernelMap(cmap, c, offs)
    t2 = time.perf_counter()
    rng = np.random.default_rng(1)
    x = rng.standard_normal((n, C_IN), dtype=np.float32)
    w = (rng.standard_normal((KS ** 3, C_IN, C_OUT), dtype=np.float32) / np.sqrt(27 * C_IN)).astype(np.float32)
    dy = rng.standard_normal((n, C_OUT), dtype=np.float32)
    CR.conv_forward(x, w, km)                      # warm
    ts = []
    for _ in range(reps):
        a = time.perf_counter()
        CR.conv_forward(x, w, km)
        CR.conv_backward(x, dy, w, km)
        ts.append(time.perf_counter() - a)
    return dict(n=n, P=int(km.counts.sum()), step_s=float(np.median(ts)), insert_s=t1 - t0, kmap_s=t2 - t1,
                cores=CR.num_threads())


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    R = 96                                          # ~125 k voxels per step: a bounded sample of the workload
    _ = cpu_sample(24, 1)                           # build/load + warm-up
    times, info = [], None
    for i in range(args.warmup + args.steps):
        info = cpu_sample(R, 1, seed=i)
        if i >= args.warmup:
            times.append(info["step_s"])
    t = float(np.mean(times))
    val = info["n"] / t
    sample = f"sphere R={R} ({info['n']} voxels, P={info['P']}), 3^3 {C_IN}->{C_OUT} fp32 fwd+bwd, kernel map cached"
    line = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": "north_star_conv_1M_k3_c64 (bounded sample per step)", "sample": sample},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": info["cores"], "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
def run_engine(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this engine has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    host_info = bind_rank_to_cpus(local, world) if world > 1 and not args.no_bind else {"cpus_total": os.cpu_count()}
    if world > 1:
        import datetime
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=180))
    import MinkowskiEngine as ME
    import MinkowskiEngineBackend._C as _C
    from mink_octtree_stablediffusion_b200 import _lib
    from mink_octtree_stablediffusion_b200.parallel import allreduce_gradients
    lib = _lib.load(build_if_missing=False)        # fail loudly if the CUDA library is missing
    dev = torch.device("cuda", local)
    pk = peaks()

    R = args.radius
    c = make_cloud(R, seed=rank)
    n = len(c)
    g = torch.Generator().manual_seed(1234 + rank)
    coords_h = torch.from_numpy(c).pin_memory()
    feats_h = torch.randn(n, C_IN, generator=g).to(torch.bfloat16).pin_memory()
    dy = torch.randn(n, C_OUT, generator=g).to(torch.bfloat16).to(dev)
    torch.manual_seed(0)
    conv = ME.MinkowskiConvolution(C_IN, C_OUT, kernel_size=KS, dimension=3).to(torch.bfloat16).to(dev)
    kg = conv.kernel_generator
    coords_d = coords_h.to(dev)
    feats_d = feats_h.to(dev)

    # resident state: coordinate map + cached kernel map
    st = ME.SparseTensor(feats_d, coords_d)
    mgr, key = st.coordinate_manager._manager, st.coordinate_map_key
    empty_off = torch.empty(0, dtype=torch.int32)
    conv_args = (kg.kernel_size, kg.kernel_stride, kg.kernel_dilation, kg.region_type, empty_off)

    pending = [None, None]      # (work handle, tensor) of the all-reduce in flight

    def step_resident(ev=None):
        if ev:
            ev[0].record()
        out = _C.ConvolutionForwardGPU(feats_d, conv.kernel, *conv_args, False, conv.convolution_mode, key, key, mgr)
        if ev:
            ev[1].record()
        gi, gw = _C.ConvolutionBackwardGPU(feats_d, dy, conv.kernel, *conv_args, conv.convolution_mode, key, key, mgr)
        if ev:
            ev[2].record()
        if world > 1:
            # weight gradients only.  Asynchronous: the collective runs on NCCL's stream while
            # the next step's forward is already on the compute stream; it is waited for one step later
            # (a training loop consumes the reduced gradient in the optimizer, after the next launch burst)
            if pending and pending[0] is not None:
                pending[0].wait()
            pending[0] = dist.all_reduce(gw, async_op=True)
            pending[1] = gw
        return out, gi, gw

    out, gi, gw = step_resident()
    torch.cuda.synchronize()
    kmap = next(iter(mgr._kmaps.values()))
    P = int((kmap.nbr[:, :n] >= 0).sum().item())      # (in, out) pairs of the map
    conv_path = int(lib.mk_conv_last_path())

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident timing ------------------------------------------
    for _ in range(args.warmup):
        step_resident()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    sync_all()
    launches0 = int(lib.mk_launch_count())
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_start.record()
    for i in range(args.steps):
        step_resident(evs[i])
    if pending[0] is not None:
        pending[0].wait()                           # the last step's all-reduce belongs to the timed region
    t_end.record()
    sync_all()
    launches = int(lib.mk_launch_count()) - launches0
    clocks = sampler.stop() if rank == 0 else None
    ms_total = t_start.elapsed_time(t_end)
    fwd_ms = float(np.mean([e[0].elapsed_time(e[1]) for e in evs]))
    bwd_ms = float(np.mean([e[1].elapsed_time(e[2]) for e in evs]))

    # ---- end to end through the public ME API, host buffers ------------------
    grad_h = torch.empty_like(conv.kernel, device="cpu").pin_memory()
    chk_h = torch.empty(1, dtype=torch.float32).pin_memory()

    # Software pipeline over two buffer sets: the host->device copy of step i+1 (copy stream) overlaps step i's
    # kernels, and the host reads step i's result while step i+1 is already queued.  Every step still copies its
    # own inputs from pinned host memory and has its result read back on the host, all inside the timed region.
    copy_stream = torch.cuda.Stream(dev)
    bufs = [(coords_d, feats_d), (torch.empty_like(coords_d), torch.empty_like(feats_d))]
    ready = [torch.cuda.Event(), torch.cuda.Event()]     # inputs of the slot have landed
    done = [torch.cuda.Event(), torch.cuda.Event()]      # the step that used the slot has finished
    grad_hs = [grad_h, torch.empty_like(grad_h).pin_memory()]
    chk_hs = [chk_h, torch.empty_like(chk_h).pin_memory()]
    used = [False, False]

    def issue_copy(slot):
        with torch.cuda.stream(copy_stream):
            if used[slot]:
                copy_stream.wait_event(done[slot])       # the previous reader of these buffers (device-side wait)
            bufs[slot][0].copy_(coords_h, non_blocking=True)
            bufs[slot][1].copy_(feats_h, non_blocking=True)
            ready[slot].record(copy_stream)

    def launch_step(slot):
        main = torch.cuda.current_stream()
        main.wait_event(ready[slot])
        conv.kernel.grad = None
        x = bufs[slot][1].detach().requires_grad_(True)   # so that backward computes the data gradient too
        s = ME.SparseTensor(x, bufs[slot][0])        # new manager: hash insert
        o = conv(s)                                  # kernel-map build + forward
        o.F.backward(dy)                             # dgrad + wgrad
        if world > 1:
            allreduce_gradients([conv.kernel])       # the only exchange of the data-parallel step
        grad_hs[slot].copy_(conv.kernel.grad, non_blocking=True)
        chk_hs[slot].copy_(o.F.detach()[:1024].float().sum().view(1), non_blocking=True)   # detached: the host
        # buffer must not keep the step's autograd graph (and its maps / tables) alive
        done[slot].record(main)
        used[slot] = True
        del s, o, x                                  # maps / tables go back to the caching allocator now

    def finish_step(slot):
        done[slot].synchronize()                     # the step's result is on the host
        return float(chk_hs[slot][0].detach())

    def run_e2e(nsteps, wall=None):
        issue_copy(0)
        for i in range(nsteps):
            t_a = time.perf_counter()
            if i + 1 < nsteps:
                issue_copy((i + 1) & 1)
            t_b = time.perf_counter()
            launch_step(i & 1)
            t_c = time.perf_counter()
            if i > 0:
                finish_step((i - 1) & 1)
            if wall is not None:
                wall.append((1e3 * (time.perf_counter() - t_a), 1e3 * (t_b - t_a), 1e3 * (t_c - t_b)))
        finish_step((nsteps - 1) & 1)

    # every step builds a new coordinate manager (64 MB table, 108 MB neighbour table, workspaces):
    # warm the caching allocator until those sizes are all served from its cache
    import gc
    gc.collect()
    recent = []
    for i in range(8):                                # >= max(W, 8) warm-up steps, until the step time settles
        t_a = time.perf_counter()
        run_e2e(4)
        torch.cuda.synchronize()
        recent.append(time.perf_counter() - t_a)
        if 4 * (i + 1) >= max(args.warmup, 8) and max(recent[-2:]) < 1.15 * min(recent[-2:]):
            break
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_wall = []
    ms0 = torch.cuda.memory_stats(dev)
    run_e2e(args.steps, e2e_wall)
    e1.record()
    if rank == 0 and os.environ.get("BENCH_VERBOSE"):
        ms1 = torch.cuda.memory_stats(dev)
        print("e2e per-step wall ms (total/copy issue/launch):",
              " ".join(f"{t[0]:.2f}/{t[1]:.2f}/{t[2]:.2f}" for t in e2e_wall), file=sys.stderr)
        print("allocator: segments %d -> %d, retries %d -> %d, reserved %.1f GB" % (
            ms0["segment.all.current"], ms1["segment.all.current"], ms0["num_alloc_retries"],
            ms1["num_alloc_retries"], ms1["reserved_bytes.all.current"] / 1e9), file=sys.stderr)
    sync_all()
    ms_e2e = e0.elapsed_time(e1)
    # the ceiling of the e2e leg: nothing but its host-to-device copies, all ranks at once
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(copy_stream):
        for i in range(3):
            bufs[i & 1][1].copy_(feats_h, non_blocking=True)
        copy_stream.synchronize()
        if world > 1:
            dist.barrier()
        h0.record(copy_stream)
        for i in range(10):
            bufs[i & 1][0].copy_(coords_h, non_blocking=True)
            bufs[i & 1][1].copy_(feats_h, non_blocking=True)
        h1.record(copy_stream)
    copy_stream.synchronize()
    ms_h2d = h0.elapsed_time(h1) / 10
    sync_all()

    # ---- integer stages, timed alone (new map each iteration) ----------------
    stage = {}
    if rank == 0:
        from mink_octtree_stablediffusion_b200 import backend_ops as _ops
        ins, kmb, kme, kto = [], [], [], []
        busy = torch.empty((8192, 8192), dtype=torch.bfloat16, device=dev).normal_()
        for _ in range(7):
            a, b, c2, d2, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(5))
            m2 = _C.CoordinateMapManagerGPU_c10()
            torch.mm(busy, busy)     # ~1 ms of device work: the host runs ahead, the events time the device only
            a.record()
            k2, _maps = m2.insert_and_map(coords_d, [1, 1, 1], "")
            b.record()
            # the table the conv kernels read (mk_kmap_build) ...
            _ops._kernel_map(m2, k2, k2, kg.kernel_size, kg.kernel_stride, kg.kernel_dilation, kg.region_type,
                             empty_off, False, False)
            c2.record()
            # ... and the (in, out) pair lists the Python API exports from it
            m2.kernel_map(k2, k2, kg.kernel_size, kg.kernel_stride, kg.kernel_dilation, kg.region_type, empty_off,
                          False, False)
            d2.record()
            # ... and the tile-ordered form the tensor-core kernels read (sort by block + neighbour pattern,
            # permuted table, tile mask); built once per kernel map, on its first 16-bit convolution
            _ops._ordered(m2, next(iter(m2._kmaps.values())))
            e2.record()
            torch.cuda.synchronize()
            ins.append(a.elapsed_time(b))
            kmb.append(b.elapsed_time(c2))
            kme.append(c2.elapsed_time(d2))
            kto.append(d2.elapsed_time(e2))
        # the insert's kernels alone (mk_map_insert2 through the C-ABI, buffers allocated once): the API-level
        # figure above also holds six allocations and the device-to-host read of the unique count
        import ctypes as _ct
        cap_ = int(lib.mk_hash_capacity(n))
        tb_ = torch.empty(cap_ * 16, dtype=torch.uint8, device=dev)
        ws_ = torch.empty(int(lib.mk_insert2_workspace_bytes(n)), dtype=torch.uint8, device=dev)
        ui_, iv_ = torch.empty(n, dtype=torch.int64, device=dev), torch.empty(n, dtype=torch.int64, device=dev)
        co_, mt_ = torch.empty_like(coords_d), torch.empty(2, dtype=torch.int64, device=dev)
        vp = lambda t_: _ct.c_void_p(t_.data_ptr())   # noqa: E731
        st_ = _ct.c_void_p(torch.cuda.current_stream().cuda_stream)

        def raw_insert(flags):
            _lib.check(lib.mk_map_insert2(vp(coords_d), n, 4, vp(tb_), cap_, flags, vp(ui_), vp(iv_), vp(co_), vp(ws_),
                                          vp(mt_), _ct.c_void_p(mt_.data_ptr() + 8), st_))
        raw = {}
        for nm, fl_ in (("hash_insert_kernels_ms", 2), ("hash_insert_unique_kernels_ms", 3)):
            raw_insert(fl_)
            e_a, e_b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.mm(busy, busy)
            e_a.record()
            for _ in range(10):
                raw_insert(fl_)
            e_b.record()
            torch.cuda.synchronize()
            raw[nm] = e_a.elapsed_time(e_b) / 10
        del tb_, ws_, ui_, iv_, co_
        K = KS ** 3
        # the coordinate stages of the octree decoder / encoder on the same cloud, through the manager the
        # ME API uses (SURVEY 8d bytes): stride-2 map, k2s2 generative children, pruning of half the rows
        from mink_octtree_stablediffusion_b200 import backend_pool as _pool
        sms_, exs_, prs_ = [], [], []
        n_half = n_str = n_chl = 0
        keep_ = (torch.arange(n, device=dev) % 2 == 0)
        feat_ = torch.zeros((n, C_IN), dtype=torch.bfloat16, device=dev)
        offs8 = np.array([[i, j, k_] for i in (0, 1) for j in (0, 1) for k_ in (0, 1)], dtype=np.int64)
        for _ in range(5):
            m3 = _C.CoordinateMapManagerGPU_c10()
            k3, _ = m3.insert_and_map(coords_d, [1, 1, 1], "")
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
            torch.mm(busy, busy)
            ev[0].record()
            ks2 = m3.stride(k3, [2, 2, 2])
            ev[1].record()
            torch.mm(busy, busy)
            ev[2].record()
            kch = m3._octree_children(ks2, offs8, [1, 1, 1])
            ev[3].record()
            torch.mm(busy, busy)
            ev[4].record()
            okey = _C.CoordinateMapKey(4)
            _pool.PruningForwardGPU(feat_, keep_, k3, okey, m3)
            ev[5].record()
            torch.cuda.synchronize()
            sms_.append(ev[0].elapsed_time(ev[1]))
            exs_.append(ev[2].elapsed_time(ev[3]))
            prs_.append(ev[4].elapsed_time(ev[5]))
            n_str, n_chl, n_half = m3.size(ks2), (m3.size(kch) if kch is not None else 0), m3.size(okey)
        del feat_, keep_
        stage = {"hash_insert_ms": float(np.median(ins)), **raw, "kmap_build_ms": float(np.median(kmb)),
                 "kmap_pairs_export_ms": float(np.median(kme)), "kmap_tile_order_ms": float(np.median(kto)),
                 "hash_insert_alg_bytes": 84 * n, "kmap_alg_bytes": int((16 + 32 * K) * n),
                 "stride_ms": float(np.median(sms_)), "stride_alg_bytes": int((16 + 84) * n + 16 * n_str),
                 "expand_ms": float(np.median(exs_)), "expand_alg_bytes": int(16 * n_str + (16 + 84) * n_chl),
                 "prune_ms": float(np.median(prs_)),
                 "prune_alg_bytes": int((1 + 16 + 2 * C_IN) * n + (16 + 2 * C_IN + 84) * n_half),
                 "stage_note": "stride / expand / prune are timed at the ME-manager level: each includes its output map's "
                               "insert, allocations and (stride, prune) one count read-back"}

    # ---- wider layers of the same cloud (MinkUNet / VAE widths): tensor-bound there ----------
    wide, cfg1 = [], None
    if rank == 0 and not args.no_extra:
        for cw in (128, 256):
            gw_ = torch.Generator().manual_seed(cw)
            xw = torch.randn(n, cw, generator=gw_).to(torch.bfloat16).to(dev)
            dyw = torch.randn(n, cw, generator=gw_).to(torch.bfloat16).to(dev)
            ww = (torch.randn(KS ** 3, cw, cw, generator=gw_) / float(np.sqrt(27 * cw))).to(torch.bfloat16).to(dev)

            def lay(ev=None):
                if ev:
                    ev[0].record()
                o_ = _C.ConvolutionForwardGPU(xw, ww, *conv_args, False, conv.convolution_mode, key, key, mgr)
                if ev:
                    ev[1].record()
                r_ = _C.ConvolutionBackwardGPU(xw, dyw, ww, *conv_args, conv.convolution_mode, key, key, mgr)
                if ev:
                    ev[2].record()
                return o_, r_
            for _ in range(3):
                lay()
            evw = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(10)]
            for e_ in evw:
                lay(e_)
            torch.cuda.synchronize()
            f_ms = float(np.mean([e_[0].elapsed_time(e_[1]) for e_ in evw]))
            b_ms = float(np.mean([e_[1].elapsed_time(e_[2]) for e_ in evw]))
            fl = 2.0 * P * cw * cw
            wide.append({"kernel": f"conv forward {cw}->{cw} (tcgen05)", "bound": "tensor",
                         "achieved": fl / (f_ms * 1e-3) / 1e12, "peak": pk["tf_burst"], "unit": "TFLOP/s",
                         "frac": fl / (f_ms * 1e-3) / 1e12 / pk["tf_burst"], "fwd_ms": f_ms, "bwd_ms": b_ms,
                         "bwd_achieved": 2 * fl / (b_ms * 1e-3) / 1e12, "traffic": None})
            del xw, dyw, ww
        # BASELINE configs[0]: batch 2 x R=61 (~100 k voxels), 32 -> 64, fp32, through the ME API, beside
        # the C restatement on the identical tensors
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from golden import configs as _cfg        # seeded inputs shared with the parity tests
        c1, x1, w1, dy1 = _cfg.inputs("cfg1")
        conv1 = ME.MinkowskiConvolution(32, 64, kernel_size=KS, dimension=3).to(dev)
        with torch.no_grad():
            conv1.kernel.copy_(torch.from_numpy(w1))
        c1d, x1d, dy1d = torch.from_numpy(c1).to(dev), torch.from_numpy(x1).to(dev), torch.from_numpy(dy1).to(dev)
        t_new, t_step = [], []
        for it in range(8):
            a_, b_, c_ = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            torch.mm(busy, busy)
            a_.record()
            s1 = ME.SparseTensor(x1d.detach().requires_grad_(True), c1d)     # hash insert
            o1 = conv1(s1)                                                    # kernel map + forward
            o1.F.backward(dy1d)
            b_.record()
            conv1.kernel.grad = None
            o1 = conv1(ME.SparseTensor(s1.F.detach().requires_grad_(True), coordinate_map_key=s1.coordinate_map_key,
                                       coordinate_manager=s1.coordinate_manager))   # maps cached
            o1.F.backward(dy1d)
            c_.record()
            torch.cuda.synchronize()
            if it >= 2:
                t_new.append(a_.elapsed_time(b_))
                t_step.append(b_.elapsed_time(c_))
        from oracle import cpu_ref as CR
        from oracle import oracle as O
        CR.set_num_threads(len(os.sched_getaffinity(0)))
        km1 = CR.KernelMap(CR.CoordMap(c1), c1, O.kernel_offsets([KS] * 3, [1, 1, 1]))
        tc_ = []
        for _ in range(3):
            t0_ = time.perf_counter()
            CR.conv_forward(x1, w1, km1)
            CR.conv_backward(x1, dy1, w1, km1)
            tc_.append(time.perf_counter() - t0_)
        cfg1 = {"workload": "BASELINE configs[0]: batch 2 x sphere R=61, 3^3, 32->64, fp32 (ME API)",
                "voxels": int(len(c1)), "gpu_fwd_bwd_ms_maps_cached": float(np.median(t_step)),
                "gpu_insert_kmap_fwd_bwd_ms": float(np.median(t_new)),
                "gpu_voxels_per_s": len(c1) / (float(np.median(t_step)) * 1e-3),
                "cpu_ref_fwd_bwd_ms": float(np.median(tc_)) * 1e3, "cpu_ref_voxels_per_s": len(c1) / float(np.median(tc_)),
                "cpu_cores": CR.num_threads(), "conv_path": "tcgen05 (bf16 x3 split)" if lib.mk_conv_last_path() else "simt"}

    # ---- the other example networks (BASELINE configs[2] and [3]), single GPU only ---------------
    nets = None
    if world == 1 and not args.no_extra and not args.no_nets:
        gc.collect()
        torch.cuda.empty_cache()
        try:
            nets = run_nets_leg(args, dev, lib)
        except Exception as e:
            nets = {"error": f"{type(e).__name__}: {e}"}

    # ---- VAE leg (the other half of BASELINE.json's metric) ---------------------
    vae = None
    if not args.no_vae:
        del st, out, gi, gw, bufs
        gc.collect()
        torch.cuda.empty_cache()
        try:
            vae = run_vae_leg(args, dist, rank, world, dev, lib, args.vae_steps, args.vae_warmup)
        except Exception as e:      # the conv line must still be printed
            vae = {"error": f"{type(e).__name__}: {e}"}
            if world > 1:
                raise

    # ---- reduce over ranks -----------------------------------------------------
    t = torch.tensor([ms_total, ms_e2e, fwd_ms, bwd_ms, ms_h2d], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e, fwd_ms, bwd_ms, ms_h2d = t.tolist()
    if rank == 0:
        ms_step = ms_total / args.steps
        value = n * world / (ms_step * 1e-3)
        e2e_val = n * world / (ms_e2e / args.steps * 1e-3)
        flops_pass = 2.0 * P * C_IN * C_OUT
        tf = flops_pass / (fwd_ms * 1e-3) / 1e12
        peak = pk["tf_burst"]      # the timed loop lasts tens of ms at full clocks: burst conditions
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "north_star_conv_1M_k3_c64", "voxels_per_gpu": n, "pairs_per_gpu": P,
                       "generator": f"sphere surface R={R}", "kernel": "3x3x3 stride 1", "channels": "64->64",
                       "step": "conv fwd + dgrad + wgrad, kernel map (table + tile-ordered form) cached", "parallelism": f"dp{world}",
                       "l2": "working set (features, grads, neighbour table ~0.7 GB) exceeds the 126 MB L2"},
            "roofline": {"bound": "tensor", "kernel": "conv forward (%s)" % ("tcgen05" if conv_path else "simt"),
                         "achieved": tf, "peak": peak, "unit": "TFLOP/s", "frac": tf / peak,
                         # dram__bytes_read+write of one k_conv_tc launch at R=271 on the tile-ordered table,
                         # ncu --set full (profiles/r2b_ncu_full_conv64.csv: 204.9 MB read + 95.9 MB written);
                         # algorithmic bytes are x + y + neighbour table + output-row permutation
                         "traffic": 300.8e6 if (conv_path and R == R_FULL) else None,
                         "traffic_source": "ncu --set full capture profiles/r2b_ncu_full_conv64.csv (same kernel, "
                                           "same cloud), not measured in this run",
                         "alg_bytes_per_launch": int(n * (C_IN + C_OUT) * 2 + 4 * KS ** 3 * n + 4 * n),
                         "peak_source": pk["source"] + " (burst bf16 matmul)",
                         "useful_flops_per_launch": flops_pass},
            "stages": {"conv_fwd_ms": fwd_ms, "conv_bwd_ms": bwd_ms,
                       "fwd_bwd_useful_tflops": 3 * flops_pass / ((fwd_ms + bwd_ms) * 1e-3) / 1e12, **stage},
            "e2e": {"value": e2e_val, "unit": UNIT,
                    "h2d_bytes_per_step": int(coords_h.numel() * 4 + feats_h.numel() * 2),
                    "d2h_bytes_per_step": int(grad_h.numel() * 2 + 4),
                    "h2d_only_ms_per_step": ms_h2d, "h2d_ceiling_voxels_per_s": n * world / (ms_h2d * 1e-3),
                    "h2d_gb_per_s_per_rank": (coords_h.numel() * 4 + feats_h.numel() * 2) / (ms_h2d * 1e-3) / 1e9,
                    "host": host_info,
                    "path": "pinned host coords+features -> ME.SparseTensor (hash insert) -> MinkowskiConvolution "
                            "(kernel-map build + fwd) -> backward -> grad_kernel + checksum to host",
                    "pipeline": "two buffer sets: step i+1's H2D copy (copy stream) and launch overlap step i's kernels "
                                "and result read-back"},
            "gpu_launches": launches, "clocks": clocks,
        }
        if stage:
            line["stages"]["hash_insert_hbm_frac"] = stage["hash_insert_alg_bytes"] / (stage["hash_insert_kernels_ms"] * 1e-3) / 1e9 / pk["hbm"]
            line["stages"]["hash_insert_api_hbm_frac"] = stage["hash_insert_alg_bytes"] / (stage["hash_insert_ms"] * 1e-3) / 1e9 / pk["hbm"]
            line["stages"]["kmap_hbm_frac"] = stage["kmap_alg_bytes"] / (stage["kmap_build_ms"] * 1e-3) / 1e9 / pk["hbm"]
            for nm in ("stride", "expand", "prune"):
                line["stages"][nm + "_hbm_frac"] = stage[nm + "_alg_bytes"] / (stage[nm + "_ms"] * 1e-3) / 1e9 / pk["hbm"]
        if world == 1 and not args.no_cpu_baseline:
            s = cpu_sample(96, 3)
            line["cpu_baseline"] = {"value": s["n"] / s["step_s"], "unit": UNIT, "cores": s["cores"], "kind": "port",
                                    "sample": f"sphere R=96 ({s['n']} voxels), 3^3 64->64 fp32 fwd+bwd, kernel map "
                                              f"cached, median of 3 (oracle/cpu_ref.c, OpenMP)"}
        if wide:
            line["roofline_wide"] = wide
        if cfg1:
            line["cfg1"] = cfg1
        if vae is not None:
            line["vae"] = vae
        if nets is not None:
            line["nets"] = nets
        emit(line)
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------
# VAE leg: the reference's unchanged octree VAE (examples/ae_res.py) on the engine
# ---------------------------------------------------------------------------
_BUCKETS = (("conv", ("k_conv_", "k_pack_b", "k_wgrad_reduce", "k_split_")),
            ("kernel_map", ("k_kmap", "k_tile_mask", "k_order_key", "k_permute_table", "k_pairs", "k_radix", "k_child")),
            ("hash_insert", ("k_insert", "k_first_count", "k_scan_block", "k_compact_first", "k_finalize",
                             "k_table_clear", "k_lookup", "k_dedup")),
            ("coords_prune", ("k_coords_", "k_flag_", "k_rows_gather", "k_expand")),
            ("norm", ("batch_norm", "k_seg_", "k_segment", "k_broadcast", "k_global")),
            ("optimizer", ("multi_tensor_apply", "Optimizer")),
            ("sort_unique", ("RadixSort", "DeviceScan", "unique", "DeviceSelect", "DeviceReduce")),
            ("copies", ("Memcpy", "Memset")))


def _bucket(name: str) -> str:
    for b, keys in _BUCKETS:
        if any(k in name for k in keys):
            return b
    return "elementwise_other"


def run_vae_leg(args, dist, rank, world, dev, lib, steps, warmup):
    """VAE steps/s: the unmodified `VaeModule.training_step` (examples/ae_res.py:817-853; VAE.forward
    :716-789) + backward + Adam on BASELINE config 2 (batch 16 per GPU @ 128^3 G3 shapes,
    vae_channel (32,128,512,512,4)).  Per step: coordinates copied from pinned host memory, loss read
    back by the step itself (`.item()`).  N > 1: the reference's DDP recipe - SyncBatchNorm
    (ae_res.py:996-997) and a gradient all-reduce; every rank holds its own 16 shapes (weak scaling)."""
    import gc
    import torch
    from harness import workloads as W
    from mink_octtree_stablediffusion_b200.parallel import allreduce_gradients
    import MinkowskiEngine as ME
    res, batch = args.vae_res, args.vae_batch
    cfg = W.vae_config(resolution=res, batch_size=batch)
    coords_h = torch.from_numpy(W.batched(W.g3_batch(res, batch, first_index=rank * batch))).pin_memory()
    n_vox = int(coords_h.shape[0])
    out = {}
    for mode in args.vae_dtypes.split(","):
        # fp32: the engine's default fp32 policy (tensor cores where 1e-5 holds, SIMT FFMA elsewhere);
        # fp32_tc: every fp32 convolution on the tensor cores (bf16 split, ~1e-4 on the 512-channel layers)
        os.environ["MK_F32_TC"] = "2" if mode == "fp32_tc" else "1"
        torch.manual_seed(0)
        model, plm, _ = W.load_vae(cfg)
        if world > 1:
            model = ME.MinkowskiSyncBatchNorm.convert_sync_batchnorm(model)
            plm.model = model
        plm = plm.to(dev)
        plm.train()
        opt = plm.configure_optimizers()
        params = [p for p in plm.parameters() if p.requires_grad]
        amp = torch.autocast("cuda", dtype=torch.bfloat16, enabled=(mode == "bf16"))

        def step():
            batch_d = {"coords": coords_h.to(dev, non_blocking=True)}
            opt.zero_grad(set_to_none=True)
            with amp:
                loss = plm.training_step(batch_d, 0)
            loss.backward()
            if world > 1:
                allreduce_gradients(params, average=True)
            opt.step()
            return loss

        for _ in range(warmup):
            loss = step()
        # the reference's training_step calls gc.collect() every step (ae_res.py:822); after warm-up the
        # long-lived objects (modules, torch, the staged scripts) are frozen out of the collector's scan,
        # so that call walks only the step's own garbage (51 ms -> ~1 ms per step)
        gc.collect()
        gc.freeze()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
        l0 = int(lib.mk_launch_count())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        walls = []
        for _ in range(steps):
            t_w = time.perf_counter()
            loss = step()
            walls.append((time.perf_counter() - t_w) * 1e3)   # (the step reads its loss back: it ends synchronised)
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        rec = {"steps_per_s": 1e3 / ms, "ms_per_step": ms, "samples_per_s": batch * world * 1e3 / ms,
               "voxels_per_s": n_vox * world * 1e3 / ms, "loss": float(loss), "engine_launches_per_step":
               (int(lib.mk_launch_count()) - l0) // steps,
               # host wall time of the individual steps on this rank: the step is host-bound, so a busy host
               # core shows up as a few slow steps; steps_per_s above is the mean over all of them
               "step_wall_ms": {"median": float(np.median(walls)), "min": float(np.min(walls)),
                                "max": float(np.max(walls))}}
        if not args.no_vae_profile:
            # one more step on EVERY rank (it holds collectives); rank 0 records its kernels
            import contextlib
            from torch.profiler import ProfilerActivity, profile as tprofile
            with (tprofile(activities=[ProfilerActivity.CUDA]) if rank == 0 else contextlib.nullcontext()) as prof:
                step()
                torch.cuda.synchronize()
        if rank == 0 and not args.no_vae_profile:
            split = {}
            for e in prof.key_averages():
                if e.device_time_total > 0:
                    b = _bucket(e.key)
                    split[b] = split.get(b, 0.0) + e.device_time_total / 1e3
            busy = sum(split.values())
            split = {k: round(v, 3) for k, v in sorted(split.items(), key=lambda kv: -kv[1])}
            split["device_idle_host_gaps"] = round(max(ms - busy, 0.0), 3)
            rec["stage_ms"] = split
        out[mode] = rec
        os.environ["MK_F32_TC"] = "1"
        del plm, model, opt, params
        torch.cuda.empty_cache()
    return {"workload": f"examples/ae_res.py VAE training_step + Adam, batch {batch}/GPU @ {res}^3 (G3 shapes), "
                        f"vae_channel (32,128,512,512,4)", "voxels_per_gpu": n_vox, "scaling": "weak",
            "h2d_bytes_per_step": int(coords_h.numel() * 4), "steps": steps, "warmup": warmup,
            "host": "gc.freeze() after warm-up (the step's own gc.collect() then scans only new objects)",
            "dtypes": {"fp32": "reference default precision (tensor cores, bf16 x3 split with the accumulation chains bounded for the 1e-5 bound; FFMA kernels where that is not possible)",
                       "fp32_tc": "fp32 tensors, every convolution on the tensor cores (bf16 x3 split, MK_F32_TC=2)",
                       "bf16": "torch.autocast(bf16) = Lightning bf16-mixed"},
            **out}


def run_nets_leg(args, dev, lib):
    """Forward + backward of the reference's unchanged MinkUNet34C (examples/minkunet.py:193; BASELINE
    configs[2]: 1 M-voxel scene, bf16) and diffusion UNet (examples/diffusion.py:89; configs[3]: stride-8
    latents of 256^3 G3 octrees, unet_channel (4, 320, 640, 960), bf16) on the engine: ms per step with
    the coordinate maps rebuilt every step (a new SparseTensor per step, as a training loop does)."""
    import torch
    import MinkowskiEngine as ME
    from harness import workloads as W
    from oracle import oracle as O

    def timed(step, warm=3, it=5):
        for _ in range(warm):
            step()
        torch.cuda.synchronize()
        l0 = int(lib.mk_launch_count())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(it):
            loss = step()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / it, float(loss), (int(lib.mk_launch_count()) - l0) // it

    out = {}
    torch.manual_seed(0)
    mu = W.load_minkunet()
    net = mu.MinkUNet34C(3, 20, D=3).to(dev).to(torch.bfloat16)
    c = torch.from_numpy(O.sphere_surface(R_FULL, 0, 0)).to(dev)
    f = torch.rand(len(c), 3, device=dev).to(torch.bfloat16)
    labels = torch.randint(0, 20, (len(c),), device=dev)
    crit = torch.nn.CrossEntropyLoss()

    def step_mu():
        net.zero_grad(set_to_none=True)
        y = net(ME.SparseTensor(f, c))
        loss = crit(y.F.float(), labels)
        loss.backward()
        return loss

    ms, loss, nl = timed(step_mu)
    out["minkunet34c"] = {"workload": "examples/minkunet.py MinkUNet34C(3, 20) fwd+bwd, bf16, sphere surface R=271",
                          "voxels": int(len(c)), "ms_per_step": ms, "voxels_per_s": len(c) / (ms * 1e-3),
                          "loss": loss, "engine_launches_per_step": nl}
    del net, c, f, labels
    torch.cuda.empty_cache()

    cfg = W.diffusion_config(resolution=256, batch_size=4)
    d = W.load_unet(cfg)
    torch.manual_seed(0)
    unet = d.UNet(cfg.unet_channel, cfg).to(dev).to(torch.bfloat16)
    lat = torch.from_numpy(O.stride_coords(W.batched(W.g3_batch(256, 4)), [8, 8, 8])).to(dev)
    xf = torch.randn(len(lat), 4, device=dev).to(torch.bfloat16)
    noise = torch.randn(len(lat), 4, device=dev).to(torch.bfloat16)
    tt = torch.randint(0, 1000, (4,), device=dev)

    def step_un():
        unet.zero_grad(set_to_none=True)
        y = unet(ME.SparseTensor(xf, lat, tensor_stride=8), tt)
        loss = torch.nn.functional.mse_loss(y.F.float(), noise.float())
        loss.backward()
        return loss

    ms, loss, nl = timed(step_un)
    out["diffusion_unet"] = {"workload": "examples/diffusion.py UNet((4, 320, 640, 960)) fwd+bwd (one DDPM training "
                                         "step's network pass), bf16, stride-8 latents of 4 G3 octrees @ 256^3",
                             "latent_voxels": int(len(lat)), "ms_per_step": ms, "steps_per_s": 1e3 / ms,
                             "loss": loss, "engine_launches_per_step": nl}
    return out


_JSON_FD = None


def emit(line) -> None:
    """Write the result line to the real stdout (see main)."""
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference"])
    ap.add_argument("--radius", type=int, default=R_FULL, help="sphere radius of the synthetic cloud (271 -> 1 M voxels)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-vae", action="store_true", help="skip the VAE steps/s leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the wide-layer rooflines and the cfg1 leg")
    ap.add_argument("--no-nets", action="store_true", help="skip the MinkUNet34C / diffusion UNet leg")
    ap.add_argument("--no-bind", action="store_true", help="do not bind the ranks to their GPUs' host cores")
    ap.add_argument("--no-vae-profile", action="store_true")
    ap.add_argument("--vae-steps", type=int, default=20)
    ap.add_argument("--vae-warmup", type=int, default=5)
    ap.add_argument("--vae-res", type=int, default=128)
    ap.add_argument("--vae-batch", type=int, default=16)
    ap.add_argument("--vae-dtypes", default="bf16,fp32_tc,fp32")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    # stdout carries exactly ONE line, the JSON: everything libraries print on the way (NCCL's version
    # banner goes to stdout) is sent to stderr by pointing fd 1 at fd 2 until the line is written
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_engine(args)


if __name__ == "__main__":
    main()
