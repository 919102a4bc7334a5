"""Normalisation layers on the engine's two fused passes (csrc/norm.cu) against
(a) NumPy for the raw C-ABI entry points, (b) torch.nn.BatchNorm1d for the batch
norms, (c) the composed global-pooling + broadcast form on the oracle-backed CPU
backend in float64 for the instance-norm family (reference:
MinkowskiNormalization.py:55-195, :198-622).  fp32 <= 1e-5, bf16 <= 1e-2, fp64 <= 1e-9."""
import copy
import ctypes as C

import numpy as np
import pytest
import torch

import abi
from oracle import me_cpu_backend as CPUB
from oracle import oracle as O

pytestmark = pytest.mark.gpu

TOL = {torch.float64: 1e-9, torch.float32: 1e-5, torch.bfloat16: 1e-2}


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp(min=1e-30))


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.bfloat16])
@pytest.mark.parametrize("c", [32, 3, 512])
def test_seg_sums_and_affine_abi(lib, dtype, c):
    from mink_octtree_stablediffusion_b200 import _lib
    n, nseg = 20011, 5
    g = torch.Generator().manual_seed(c)
    a = (torch.randn(n, c, generator=g, dtype=torch.float64) + 3).to(dtype).cuda()
    b = torch.randn(n, c, generator=g, dtype=torch.float64).to(dtype).cuda()
    seg = torch.sort(torch.randint(0, nseg, (n,), generator=g))[0].int()
    seg[::97] = -1                                            # ignored rows
    seg = seg.cuda()
    acc = torch.float64 if dtype == torch.float64 else torch.float32
    sa = torch.randn(nseg, c, generator=g).to(acc).cuda()
    sb = torch.randn(nseg, c, generator=g).to(acc).cuda()
    out = torch.full((nseg, 2, c), 7.0, dtype=acc, device="cuda")
    _lib.check(lib.mk_seg_sums(abi.DT[dtype], abi.P(a), abi.P(b), abi.P(sa), abi.P(sb), abi.P(seg), n, c, nseg,
                               abi.P(out), abi.S()))
    ad, bd, s = a.double().cpu(), b.double().cpu(), seg.cpu().long()
    want = torch.zeros(nseg, 2, c, dtype=torch.float64)
    ok = s >= 0
    x0 = ad[ok] - sa.double().cpu()[s[ok]]
    want[:, 0].index_add_(0, s[ok], x0)
    want[:, 1].index_add_(0, s[ok], x0 * (bd[ok] - sb.double().cpu()[s[ok]]))
    # fp32 accumulation of ~20 k un-centred terms (the layers pass a shift near the mean)
    assert rel(out, want) <= (1e-12 if dtype == torch.float64 else 2e-4)
    ca, cb, cc = (torch.randn(nseg, c, generator=g).to(acc).cuda() for _ in range(3))
    y = torch.empty_like(a)
    _lib.check(lib.mk_seg_affine(abi.DT[dtype], abi.P(a), abi.P(b), abi.P(seg), n, c, abi.P(ca), abi.P(cb), abi.P(cc),
                                 abi.P(y), abi.S()))
    sc = s.clamp(min=0)
    wy = ca.double().cpu()[sc] * ad + cb.double().cpu()[sc] * bd + cc.double().cpu()[sc]
    wy[~ok] = 0
    assert rel(y, wy) <= TOL[dtype]


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.bfloat16])
@pytest.mark.parametrize("c", [32, 4, 128])
def test_batch_norm_matches_torch(dtype, c):
    import MinkowskiEngine as ME
    coords = torch.from_numpy(np.concatenate([O.sphere_surface(14, 0, 0), O.sphere_surface(11, 1, 1)])).cuda()
    n = len(coords)
    torch.manual_seed(c)
    x64 = torch.randn(n, c, dtype=torch.float64) * 2 + 5           # mean far from zero: the shifted sums must cope
    dy64 = torch.randn(n, c, dtype=torch.float64)
    ref = torch.nn.BatchNorm1d(c).double()
    with torch.no_grad():
        ref.weight.uniform_(0.5, 1.5)
        ref.bias.uniform_(-1, 1)
    bn = ME.MinkowskiBatchNorm(c).to(dtype if dtype != torch.bfloat16 else torch.float32)   # bf16 features, fp32
    bn.bn.load_state_dict(ref.state_dict())                                               # parameters (autocast)
    ref.load_state_dict(bn.bn.state_dict())       # both sides hold the parameter values rounded to the module dtype
    bn = bn.cuda()
    xin = x64.to(dtype).cuda().requires_grad_(True)
    xr = xin.detach().double().cpu().requires_grad_(True)
    for _ in range(2):                                             # twice: running statistics accumulate
        out = bn(ME.SparseTensor(xin, coords))
        yr = ref(xr)
    assert rel(out.F, yr) <= TOL[dtype]
    out.F.backward(dy64.to(dtype).cuda())
    yr.backward(dy64.to(dtype).double())
    tol = TOL[dtype]
    assert rel(xin.grad, xr.grad) <= (tol if dtype != torch.bfloat16 else 2e-2)
    assert rel(bn.bn.weight.grad, ref.weight.grad) <= max(tol, 1e-5) * (1 if dtype != torch.bfloat16 else 3)
    assert rel(bn.bn.bias.grad, ref.bias.grad) <= max(tol, 1e-5)
    assert rel(bn.bn.running_mean, ref.running_mean) <= max(tol, 1e-6)
    assert rel(bn.bn.running_var, ref.running_var) <= max(tol, 1e-5) * 3
    assert int(bn.bn.num_batches_tracked) == 2
    bn.eval()                                                      # eval: torch's own path on the running statistics
    ref.eval()
    assert rel(bn(ME.SparseTensor(xin.detach(), coords)).F, ref(xr.detach())) <= TOL[dtype] * 3


def test_batch_norm_of_a_16_bit_module():
    """net.to(torch.bfloat16) (how scripts/run_models.py and the MinkUNet bench leg run): parameters and
    running statistics are bf16 too; the fused passes must serve them (fp32 algebra, statistics written
    back in bf16) and agree with an fp64 BatchNorm1d on the same rounded values."""
    import MinkowskiEngine as ME
    from mink_octtree_stablediffusion_b200 import backend_norm
    coords = torch.from_numpy(O.sphere_surface(14, 0, 0)).cuda()
    n, c = len(coords), 64
    torch.manual_seed(5)
    bn = ME.MinkowskiBatchNorm(c)
    with torch.no_grad():
        bn.bn.weight.uniform_(0.5, 1.5)
        bn.bn.bias.uniform_(-1, 1)
    bn = bn.to(torch.bfloat16).cuda()
    ref = torch.nn.BatchNorm1d(c).double()
    ref.load_state_dict({k: v.double().cpu() if v.is_floating_point() else v.cpu() for k, v in bn.bn.state_dict().items()})
    x = (torch.randn(n, c) * 2 + 3).bfloat16().cuda().requires_grad_(True)
    xr = x.detach().double().cpu().requires_grad_(True)
    dy = torch.randn(n, c).bfloat16()
    calls = []
    orig = backend_norm.batch_norm_train
    backend_norm.batch_norm_train = lambda *a, **k: (calls.append(1), orig(*a, **k))[1]
    try:
        out = bn(ME.SparseTensor(x, coords))
    finally:
        backend_norm.batch_norm_train = orig
    assert calls, "the fused path must serve 16-bit parameters"
    yr = ref(xr)
    assert out.F.dtype == torch.bfloat16 and rel(out.F, yr) <= 1e-2
    out.F.backward(dy.cuda())
    yr.backward(dy.double())
    assert rel(x.grad, xr.grad) <= 2e-2
    assert bn.bn.weight.grad.dtype == torch.bfloat16 and rel(bn.bn.weight.grad, ref.weight.grad) <= 1e-2
    assert rel(bn.bn.bias.grad, ref.bias.grad) <= 1e-2
    assert bn.bn.running_mean.dtype == torch.bfloat16 and rel(bn.bn.running_mean, ref.running_mean) <= 1e-2
    assert rel(bn.bn.running_var, ref.running_var) <= 1e-2


def _norm_case(make, with_emb=False, c=32):
    """module on the GPU (fused passes) vs the same module on CPU tensors over the oracle backend (composed)."""
    import MinkowskiEngine as ME
    coords = torch.from_numpy(np.concatenate([O.sphere_surface(12, 0, 0), O.sphere_surface(9, 1, 1),
                                              O.sphere_surface(7, 2, 2)]))
    n = len(coords)
    torch.manual_seed(1)
    x = torch.randn(n, c, dtype=torch.float64) * 1.5 + 2
    dy = torch.randn(n, c, dtype=torch.float64)
    emb = torch.randn(3, 16, dtype=torch.float64)
    mod = make().double()
    with torch.no_grad():
        mod.weight.uniform_(0.5, 1.5)
        mod.bias.uniform_(-1, 1)
        if with_emb:
            mod.linear.weight.normal_(0, 0.3)
    res = {}
    for dev, dtype in (("cpu", torch.float64), ("cuda", torch.float64), ("cuda", torch.float32)):
        m = copy.deepcopy(mod).to(dtype).to(dev)
        xi = x.detach().clone().to(dtype).to(dev).requires_grad_(True)
        e = emb.detach().clone().to(dtype).to(dev).requires_grad_(True)
        ctx = CPUB.installed() if dev == "cpu" else torch.enable_grad()
        with ctx:
            st = ME.SparseTensor(xi, coords.to(dev))
            out = m(st, e) if with_emb else m(st)
            out.F.backward(dy.to(dtype).to(dev))
        res[(dev, dtype)] = (out.F, xi.grad, m.weight.grad, m.bias.grad, e.grad if with_emb else None)
    ref = res[("cpu", torch.float64)]
    for key, tol in ((("cuda", torch.float64), 1e-9), (("cuda", torch.float32), 2e-5)):
        for got, want in zip(res[key], ref):
            if want is not None:
                assert rel(got, want) <= tol, key


def test_instance_norm_family_matches_composed_oracle_path():
    import MinkowskiEngine as ME
    _norm_case(lambda: ME.MinkowskiInstanceNorm(32))
    _norm_case(lambda: ME.MinkowskiStableInstanceNorm(32, group=1))
    _norm_case(lambda: ME.MinkowskiStableInstanceNorm(32, group=4))
    _norm_case(lambda: ME.MinkowskiStableGroupNorm(32))
    _norm_case(lambda: ME.AdaStableInstanceNorm(32, 16), with_emb=True)
    _norm_case(lambda: ME.MinkowskiStableInstanceNorm(6, group=3), c=6)       # scalar-path channel count
