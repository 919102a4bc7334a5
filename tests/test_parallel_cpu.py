"""N > 1 host logic on CPU (gloo, world_size 2): samples are sharded, every rank
runs the conv fwd+bwd of ITS samples only (here: the CPU oracle stands in for
the CUDA path), and the only exchange is the weight-gradient all-reduce.  The
result must equal the single-process gradient over the whole batch."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mink_octtree_stablediffusion_b200.parallel import allreduce_gradients, shard_samples  # noqa: E402


def test_shard_samples_partition_and_balance():
    sizes = [50, 10, 10, 10, 40, 30, 5, 5]
    for world in (1, 2, 3, 4, 8):
        shards = shard_samples(sizes, world)
        assert sorted(i for s in shards for i in s) == list(range(len(sizes)))   # disjoint cover
        loads = [sum(sizes[i] for i in s) for s in shards]
        assert max(loads) - min(loads) <= max(sizes)
    assert shard_samples(sizes, 2) == shard_samples(sizes, 2)                     # deterministic
    assert shard_samples([], 2) == [[], []]


def _sample(i):
    from oracle import oracle as O
    c = O.sphere_surface(6 + i, seed=i)
    rng = np.random.default_rng(100 + i)
    x = rng.standard_normal((len(c), 4))
    dy = rng.standard_normal((len(c), 8))
    return c, x, dy


def _local_grad(samples, w):
    from oracle import oracle as O
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    gw = np.zeros_like(w)
    for i in samples:
        c, x, dy = _sample(i)
        km = O.kernel_map(c, c, offs)
        _, g = O.conv_backward(x, dy, w, km)
        gw += g
    return gw


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle as O
        sizes = [len(_sample(i)[0]) for i in range(5)]
        mine = shard_samples(sizes, world)[rank]
        w = np.random.default_rng(7).standard_normal((27, 4, 8))
        kernel = torch.nn.Parameter(torch.from_numpy(w.copy()))
        bias_free = torch.nn.Parameter(torch.zeros(3, dtype=torch.float32))     # a second dtype bucket
        kernel.grad = torch.from_numpy(_local_grad(mine, w))
        bias_free.grad = torch.full((3,), float(rank + 1))
        allreduce_gradients([kernel, bias_free])
        ret[rank] = (kernel.grad.numpy().copy(), bias_free.grad.numpy().copy(), mine)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_world2_gradient_allreduce_equals_single_process():
    world = 2
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0
    w = np.random.default_rng(7).standard_normal((27, 4, 8))
    full = _local_grad(range(5), w)
    g0, b0, s0 = ret[0]
    g1, b1, s1 = ret[1]
    assert sorted(s0 + s1) == list(range(5)) and s0 and s1
    assert np.array_equal(g0, g1)                       # every rank ends with the same gradient
    assert np.allclose(g0, full, rtol=1e-12, atol=1e-12)
    assert np.array_equal(b0, np.full(3, 3.0, np.float32)) and np.array_equal(b1, b0)
