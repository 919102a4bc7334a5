"""ME Python API on the engine (GPU): SparseTensor, CoordinateManager,
MinkowskiConvolution / ConvolutionTranspose / GenerativeConvolutionTranspose,
MinkowskiPruning, pooling; float64 gradcheck as utils/gradcheck.py:34-57."""
import numpy as np
import pytest
import torch

from conftest import random_cloud
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _me():
    import MinkowskiEngine as ME
    return ME


def test_sparse_tensor_dedup_and_properties():
    ME = _me()
    c = random_cloud(4000, 5, 3, seed=21, dup_frac=0.4)
    f = torch.arange(len(c), dtype=torch.float32).view(-1, 1).repeat(1, 2)
    st = ME.SparseTensor(f.cuda(), torch.from_numpy(c).cuda())
    uc, ui, inv = O.insert_and_map(c)
    assert np.array_equal(st.C.cpu().numpy(), uc)
    assert np.array_equal(st.F[:, 0].cpu().numpy(), ui.astype(np.float32))   # RANDOM_SUBSAMPLE = first occurrence
    assert st.tensor_stride == [1, 1, 1] and st.D == 3
    cm = st.coordinate_manager
    key, (um, im) = cm.insert_and_map(torch.from_numpy(c).cuda(), [1, 1, 1], "again")
    assert np.array_equal(um.cpu().numpy(), ui) and np.array_equal(im.cpu().numpy(), inv)
    ct = torch.from_numpy(c).cuda()
    assert torch.equal(ct[um], cm.get_coordinates(key)) and torch.equal(ct, ct[um][im])


def test_cpu_tensors_are_refused():
    ME = _me()
    with pytest.raises(RuntimeError):
        ME.SparseTensor(torch.ones(3, 1), torch.zeros(3, 4, dtype=torch.int32))


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 1e-2)])
def test_minkowski_convolution_forward_backward(dtype, tol):
    ME = _me()
    c = np.concatenate([O.sphere_surface(20, 0, 0), O.sphere_surface(17, 1, 1)])
    n = len(c)
    torch.manual_seed(0)
    conv = ME.MinkowskiConvolution(32, 64, kernel_size=3, dimension=3, bias=True).to(dtype).cuda()
    x = torch.randn(n, 32).to(dtype).cuda().requires_grad_(True)
    st = ME.SparseTensor(x, torch.from_numpy(c).cuda())
    out = st_out = conv(st)
    assert out.coordinate_map_key == st.coordinate_map_key        # stride 1: same map
    dy = torch.randn(n, 64).to(dtype).cuda()
    out.F.backward(dy)
    km = O.kernel_map(c, c, O.kernel_offsets([3, 3, 3], [1, 1, 1]))
    xr, wr = x.detach().double().cpu().numpy(), conv.kernel.detach().double().cpu().numpy()
    y_ref = O.conv_forward(xr, wr, km, n) + conv.bias.detach().double().cpu().numpy()
    gi_ref, gw_ref = O.conv_backward(xr, dy.double().cpu().numpy(), wr, km)
    rel = lambda a, b: np.abs(a.double().cpu().numpy() - b).max() / np.abs(b).max()
    assert rel(st_out.F.detach(), y_ref) <= tol
    assert rel(x.grad, gi_ref) <= tol
    assert rel(conv.kernel.grad, gw_ref) <= tol
    # manager.kernel_map export: dict{k: [2, n]} with [0]=in rows, [1]=out rows
    kmap = st.coordinate_manager.kernel_map(st.coordinate_map_key, st.coordinate_map_key, kernel_size=3)
    for k, (ir, orr) in enumerate(km):
        got = kmap[k].cpu().numpy()
        order = np.argsort(got[1], kind="stable")
        assert np.array_equal(got[0][order], ir) and np.array_equal(got[1][order], orr)


def test_strided_conv_and_transpose_roundtrip_keys():
    ME = _me()
    c = O.sphere_surface(24, 3)
    torch.manual_seed(1)
    st = ME.SparseTensor(torch.randn(len(c), 8).cuda(), torch.from_numpy(c).cuda())
    down = ME.MinkowskiConvolution(8, 16, kernel_size=3, stride=2, dimension=3).cuda()
    up = ME.MinkowskiConvolutionTranspose(16, 8, kernel_size=2, stride=2, dimension=3).cuda()
    d = down(st)
    assert d.tensor_stride == [2, 2, 2]
    assert np.array_equal(O.canonical_coords(d.C.cpu().numpy()), O.canonical_coords(O.stride_coords(c, [2, 2, 2])))
    u = up(d)
    assert u.tensor_stride == [1, 1, 1]
    assert u.coordinate_map_key == st.coordinate_map_key          # non-generative: cached finer map reused
    # value check of the transposed conv against the oracle
    dc = d.C.cpu().numpy()
    offs = O.kernel_offsets([2, 2, 2], [1, 1, 1])
    km = O.kernel_map_transpose(dc, c, offs)
    ref = O.conv_forward(d.F.detach().double().cpu().numpy(), up.kernel.detach().double().cpu().numpy(), km, len(c))
    assert np.abs(u.F.detach().double().cpu().numpy() - ref).max() <= 1e-5 * np.abs(ref).max()


def test_tile_ordered_strided_and_transposed_maps(monkeypatch):
    """The tile-ordered table on maps that are not symmetric (stride 2 down, k2 s2 transposed
    up, and their data gradients through the ordered INVERSE table): a bf16 encoder/decoder pair
    through the ME API with the ordering forced on (MK_TILE_ORDER_MIN = 0) must give what it
    gives with the ordering off - features and input gradient bit-identical (a row's sum does
    not depend on its tile), weight gradients within the bf16 tolerance."""
    ME = _me()
    c = O.sphere_surface(40, 5)
    torch.manual_seed(4)
    f0 = torch.randn(len(c), 32).bfloat16()
    w_down = (torch.randn(27, 32, 64) / 30).bfloat16()
    w_up = (torch.randn(8, 64, 32) / 20).bfloat16()

    def run(order_on):
        monkeypatch.setenv("MK_TILE_ORDER", "1" if order_on else "0")
        monkeypatch.setenv("MK_TILE_ORDER_MIN", "0")
        f = f0.clone().cuda().requires_grad_(True)
        st = ME.SparseTensor(f, torch.from_numpy(c).cuda())
        down = ME.MinkowskiConvolution(32, 64, kernel_size=3, stride=2, dimension=3).cuda().to(torch.bfloat16)
        up = ME.MinkowskiConvolutionTranspose(64, 32, kernel_size=2, stride=2, dimension=3).cuda().to(torch.bfloat16)
        with torch.no_grad():
            down.kernel.copy_(w_down.cuda())
            up.kernel.copy_(w_up.cuda())
        u = up(down(st))
        g = torch.Generator().manual_seed(9)
        dy = torch.randn(u.F.shape, generator=g).bfloat16().cuda()
        (u.F * dy).sum().backward()
        return u.F.detach(), f.grad.detach(), down.kernel.grad.detach(), up.kernel.grad.detach()

    off = run(False)
    on = run(True)
    assert torch.equal(on[0], off[0]) and torch.equal(on[1], off[1])
    for a, b in ((on[2], off[2]), (on[3], off[3])):
        assert float((a.float() - b.float()).abs().max() / b.float().abs().max()) <= 1e-2


def test_generative_transpose_then_pruning_octree_step():
    """One octree decoder level (modules/vae_block.py:279-296, examples/ae_res.py:556-609)."""
    ME = _me()
    fine_gt = O.sphere_surface(16, 5)
    coarse = O.stride_coords(fine_gt, [2, 2, 2])
    torch.manual_seed(2)
    gt = ME.SparseTensor(torch.ones(len(fine_gt), 1).cuda(), torch.from_numpy(fine_gt).cuda())
    cm = gt.coordinate_manager
    x = torch.randn(len(coarse), 16).cuda().requires_grad_(True)
    st = ME.SparseTensor(x, torch.from_numpy(coarse).cuda(), tensor_stride=2, coordinate_manager=cm)
    up = ME.MinkowskiGenerativeConvolutionTranspose(16, 8, kernel_size=2, stride=2, dimension=3).cuda()
    out = up(st)
    assert out.tensor_stride == [1, 1, 1] and len(out) == 8 * len(coarse)
    offs = O.kernel_offsets([2, 2, 2], [1, 1, 1])
    assert np.array_equal(O.canonical_coords(out.C.cpu().numpy()),
                          O.canonical_coords(O.generative_coords(coarse, offs)))
    # get_target: k=1 kernel map = membership of generated coords in the GT set
    kmap = cm.kernel_map(out.coordinate_map_key, gt.coordinate_map_key, kernel_size=1, region_type=1)
    target = torch.zeros(len(out), dtype=torch.bool, device="cuda")
    for k, m in kmap.items():
        target[m[0].long()] = True
    oc = out.C.cpu().numpy()
    gtset = set(map(tuple, fine_gt))
    assert target.cpu().numpy().tolist() == [tuple(r) in gtset for r in oc]
    assert int(target.sum()) == len(fine_gt)
    pruned = ME.MinkowskiPruning()(out, target)
    assert np.array_equal(pruned.C.cpu().numpy(), oc[target.cpu().numpy()])          # order preserved
    assert torch.equal(pruned.F, out.F[target])
    pruned.F.sum().backward()
    assert x.grad is not None and torch.isfinite(x.grad).all()
    # grad of sum through prune+transpose conv: each coarse row gets sum over kept children of W_k row sums
    kmt = O.kernel_map_transpose(coarse, oc, offs)
    w = up.kernel.detach().double().cpu().numpy()
    ref = np.zeros((len(coarse), 16))
    tnp = target.cpu().numpy()
    for k, (ir, orr) in enumerate(kmt):
        sel = tnp[orr]
        np.add.at(ref, ir[sel], w[k].sum(1)[None, :].repeat(sel.sum(), 0))
    assert np.abs(x.grad.double().cpu().numpy() - ref).max() <= 1e-5 * max(np.abs(ref).max(), 1)


def test_gradcheck_float64_conv_and_transpose():
    ME = _me()
    from MinkowskiEngine.MinkowskiConvolution import MinkowskiConvolutionFunction, MinkowskiConvolutionTransposeFunction
    c = np.unique(random_cloud(40, 3, 2, seed=22), axis=0)
    x = torch.randn(len(c), 3, dtype=torch.float64).cuda().requires_grad_(True)
    st = ME.SparseTensor(x, torch.from_numpy(c).cuda())
    conv = ME.MinkowskiConvolution(3, 2, kernel_size=3, stride=2, dimension=3).double().cuda()
    cm = st.coordinate_manager
    out_key = cm.stride(st.coordinate_map_key, [2, 2, 2])
    assert torch.autograd.gradcheck(
        MinkowskiConvolutionFunction.apply,
        (x, conv.kernel, conv.kernel_generator, conv.convolution_mode, st.coordinate_map_key, out_key, cm),
        eps=1e-6, atol=1e-5, rtol=1e-3)
    xs = torch.randn(cm.size(out_key), 2, dtype=torch.float64).cuda().requires_grad_(True)
    tr = ME.MinkowskiConvolutionTranspose(2, 3, kernel_size=2, stride=2, dimension=3).double().cuda()
    assert torch.autograd.gradcheck(
        MinkowskiConvolutionTransposeFunction.apply,
        (xs, tr.kernel, tr.kernel_generator, tr.convolution_mode, out_key, st.coordinate_map_key, cm),
        eps=1e-6, atol=1e-5, rtol=1e-3)


@pytest.mark.parametrize("mode", ["max", "avg", "sum"])
def test_local_pooling_matches_oracle(mode):
    ME = _me()
    c = O.sphere_surface(14, 6)
    torch.manual_seed(3)
    x = torch.randn(len(c), 6).cuda().requires_grad_(True)
    st = ME.SparseTensor(x, torch.from_numpy(c).cuda())
    pool = {"max": ME.MinkowskiMaxPooling, "avg": ME.MinkowskiAvgPooling, "sum": ME.MinkowskiSumPooling}[mode](
        kernel_size=2, stride=2, dimension=3)
    out = pool(st)
    oc = out.C.cpu().numpy()
    km = O.kernel_map(c, oc, O.kernel_offsets([2, 2, 2], [1, 1, 1]))
    ref, _ = O.pool_forward(x.detach().double().cpu().numpy(), km, len(oc), mode)
    assert np.abs(out.F.detach().double().cpu().numpy() - ref).max() <= 1e-5
    out.F.sum().backward()
    assert torch.isfinite(x.grad).all()
    if mode == "sum":
        assert torch.allclose(x.grad, torch.ones_like(x.grad))


def test_global_pooling_and_broadcast():
    ME = _me()
    c = np.concatenate([O.sphere_surface(8, 0, 0), O.sphere_surface(9, 1, 1), O.sphere_surface(7, 2, 2)])
    torch.manual_seed(4)
    x = torch.randn(len(c), 5).cuda()
    st = ME.SparseTensor(x, torch.from_numpy(c).cuda())
    g = ME.MinkowskiGlobalAvgPooling()(st)
    ref, cnt = O.global_pool(x.double().cpu().numpy(), c[:, 0], 3, "avg")
    assert np.abs(g.F.double().cpu().numpy() - ref).max() <= 1e-5
    b = ME.MinkowskiBroadcastAddition()(st, g)
    assert np.abs(b.F.double().cpu().numpy() - (x.double().cpu().numpy() + ref[c[:, 0]])).max() <= 1e-5
    m = ME.MinkowskiBroadcastMultiplication()(st, g)
    assert np.abs(m.F.double().cpu().numpy() - (x.double().cpu().numpy() * ref[c[:, 0]])).max() <= 1e-5


def test_sparse_tensor_arithmetic_union():
    ME = _me()
    a = np.unique(random_cloud(300, 4, 1, seed=23), axis=0)
    b = np.unique(random_cloud(300, 4, 1, seed=24), axis=0)
    A = ME.SparseTensor(torch.ones(len(a), 2).cuda(), torch.from_numpy(a).cuda())
    B = ME.SparseTensor(2 * torch.ones(len(b), 2).cuda(), torch.from_numpy(b).cuda(),
                        coordinate_manager=A.coordinate_manager)
    Ssum = A + B
    sa, sb = set(map(tuple, a)), set(map(tuple, b))
    assert len(Ssum) == len(sa | sb)
    vals = {tuple(r): float(v) for r, v in zip(Ssum.C.cpu().numpy(), Ssum.F[:, 0].cpu().numpy())}
    for r in sa | sb:
        assert vals[r] == (1.0 if r in sa else 0.0) + (2.0 if r in sb else 0.0)


def test_empty_and_single_voxel_tensors():
    """Edge sizes through the ME API: pruning that keeps nothing gives an empty tensor that convolutions,
    norms' parents and the backward pass accept (zero rows in, zero rows out, zero weight gradient), and a
    single voxel convolves with its centre offset only."""
    ME = _me()
    c = torch.from_numpy(O.sphere_surface(9, 0, 0)).cuda()
    n = len(c)
    torch.manual_seed(3)
    for dtype in (torch.float32, torch.bfloat16):
        x = torch.randn(n, 16).to(dtype).cuda().requires_grad_(True)
        st = ME.SparseTensor(x, c)
        conv = ME.MinkowskiConvolution(16, 32, kernel_size=3, dimension=3).to(dtype).cuda()
        pruned = ME.MinkowskiPruning()(st, torch.zeros(n, dtype=torch.bool, device="cuda"))
        assert pruned.F.shape == (0, 16) and pruned.C.shape == (0, 4)
        out = conv(pruned)
        assert out.F.shape == (0, 32)
        out.F.sum().backward()
        assert conv.kernel.grad is not None and float(conv.kernel.grad.abs().max()) == 0.0
        assert x.grad is not None and float(x.grad.abs().max()) == 0.0
        # one voxel: only the centre offset has a neighbour
        one = ME.SparseTensor(torch.ones(1, 16).to(dtype).cuda(), c[:1].contiguous())
        y = conv(one)
        want = conv.kernel.detach()[13].double().sum(dim=0)
        assert y.F.shape == (1, 32)
        assert float((y.F.detach().double()[0] - want).abs().max()) <= (1e-5 if dtype == torch.float32 else 5e-2)
