"""GPU parity of the convolution kernels (through the C-ABI) against the CPU
oracle.  fp32: <= 1e-5 relative; bf16: <= 1e-2 relative (north_star tolerances);
fp64 checked to 1e-10."""
import numpy as np
import pytest
import torch

import abi
from conftest import random_cloud
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def _setup(coords, out_coords, offs):
    tin = abi.insert(coords)
    nbr, ld, _ = abi.kmap_build(torch.from_numpy(out_coords).cuda(), tin["table"], tin["cap"], offs)
    return nbr, ld, O.kernel_map(coords, out_coords, offs)


TOL = {torch.float64: 1e-10, torch.float32: 1e-5, torch.bfloat16: 1e-2, torch.float16: 1e-2}


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("cin,cout", [(32, 64), (64, 64), (3, 32), (1, 4), (96, 48), (128, 128)])
def test_conv_fwd_bwd_same_map(dtype, cin, cout):
    c = O.sphere_surface(18, seed=1)
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    nbr, ld, ref = _setup(c, c, offs)
    n = len(c)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(n, cin, generator=g, dtype=torch.float64)
    w = torch.randn(27, cin, cout, generator=g, dtype=torch.float64) / np.sqrt(27 * cin)
    dy = torch.randn(n, cout, generator=g, dtype=torch.float64)
    xd, wd, dyd = (t.to(dtype).cuda() for t in (x, w, dy))
    # oracle on the values the kernel actually sees (after rounding to dtype)
    xr, wr, dyr = (t.double().cpu().numpy() for t in (xd, wd, dyd))
    y_ref = O.conv_forward(xr, wr, ref, n)
    gi_ref, gw_ref = O.conv_backward(xr, dyr, wr, ref)
    tol = TOL[dtype]
    y = abi.conv_fwd(xd, wd, nbr, ld, n)
    assert rel_err(y.double().cpu().numpy(), y_ref) <= tol
    # fp32 runs on the tensor cores too (bf16 operand splitting, channels padded to 16: conv_f32.cu): the
    # offsets are split over grid.z so that no accumulator chains more than ~400 MMAs (tcgen05 accumulates
    # with truncation) and the partial outputs are added in fp32 - the 1e-5 bound below checks exactly that
    if abi.tc_eligible(dtype, n, 27, cin, cout):
        assert abi.last_path() == 1
    assert abi.tc_eligible(torch.float32, n, 27, cin, cout)
    if dtype == torch.float32:
        y_simt = abi.conv_fwd(xd, wd, nbr, ld, n, legacy=True)      # the SIMT kernel still serves the old entry point
        assert abi.last_path() == 0 and rel_err(y_simt.double().cpu().numpy(), y_ref) <= tol
    # data gradient: symmetric same-map shortcut (table itself, offsets mirrored) ...
    gi = abi.conv_fwd(dyd, wd, nbr, ld, n, transpose_w=1, kflip=1)
    assert rel_err(gi.double().cpu().numpy(), gi_ref) <= tol
    # ... and through the explicit inverse table
    inv, ld_in = abi.kmap_invert(nbr, n, ld, 27, n)
    gi2 = abi.conv_fwd(dyd, wd, inv, ld_in, n, transpose_w=1, kflip=0)
    assert rel_err(gi2.double().cpu().numpy(), gi_ref) <= tol
    gw = abi.conv_wgrad(xd, dyd, nbr, ld, n, 27)
    assert rel_err(gw.double().cpu().numpy(), gw_ref) <= tol
    if dtype in (torch.bfloat16, torch.float16):
        # fp32 weight gradient of 16-bit features (fp32 parameters): the un-rounded sum
        gw32 = abi.conv_wgrad(xd, dyd, nbr, ld, n, 27, dw_f32=True)
        assert gw32.dtype == torch.float32 and rel_err(gw32.double().cpu().numpy(), gw_ref) <= 1e-5
        assert torch.equal(gw32.to(dtype), gw) or rel_err(gw32.to(dtype).double().cpu().numpy(), gw.double().cpu().numpy()) <= 1e-2


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_conv_strided_and_transposed(dtype):
    c = O.sphere_surface(22, seed=2)
    s = O.stride_coords(c, [2, 2, 2])
    tol = TOL[dtype]
    g = torch.Generator().manual_seed(1)
    # stride-2 3^3 conv: fine -> coarse
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    nbr, ld, ref = _setup(c, s, offs)
    x = torch.randn(len(c), 64, generator=g).to(dtype).cuda()
    w = (torch.randn(27, 64, 128, generator=g) / 40).to(dtype).cuda()
    y = abi.conv_fwd(x, w, nbr, ld, len(s))
    y_ref = O.conv_forward(x.double().cpu().numpy(), w.double().cpu().numpy(), ref, len(s))
    assert rel_err(y.double().cpu().numpy(), y_ref) <= tol
    dy = torch.randn(len(s), 128, generator=g).to(dtype).cuda()
    gi_ref, gw_ref = O.conv_backward(x.double().cpu().numpy(), dy.double().cpu().numpy(), w.double().cpu().numpy(), ref)
    inv, ld_in = abi.kmap_invert(nbr, len(s), ld, 27, len(c))
    gi = abi.conv_fwd(dy, w, inv, ld_in, len(c), transpose_w=1)
    assert rel_err(gi.double().cpu().numpy(), gi_ref) <= tol
    gw = abi.conv_wgrad(x, dy, nbr, ld, len(s), 27)
    assert rel_err(gw.double().cpu().numpy(), gw_ref) <= tol
    # generative k2 s2 transpose: coarse -> 8 children
    offs2 = O.kernel_offsets([2, 2, 2], [1, 1, 1])
    fine = O.generative_coords(s, offs2)
    tin = abi.insert(s)
    nbr2, ld2, _ = abi.kmap_build(torch.from_numpy(fine).cuda(), tin["table"], tin["cap"], -offs2)
    ref2 = O.kernel_map_transpose(s, fine, offs2)
    xs = torch.randn(len(s), 128, generator=g).to(dtype).cuda()
    w2 = (torch.randn(8, 128, 64, generator=g) / 12).to(dtype).cuda()
    y2 = abi.conv_fwd(xs, w2, nbr2, ld2, len(fine))
    y2_ref = O.conv_forward(xs.double().cpu().numpy(), w2.double().cpu().numpy(), ref2, len(fine))
    assert rel_err(y2.double().cpu().numpy(), y2_ref) <= tol


def test_conv_dense_grid_equals_conv3d():
    import torch.nn.functional as F
    S, cin, cout = 8, 16, 32
    gidx = np.stack(np.meshgrid(np.arange(S), np.arange(S), np.arange(S), indexing="ij"), -1).reshape(-1, 3)
    coords = np.concatenate([np.zeros((len(gidx), 1), np.int64), gidx], 1).astype(np.int32)
    coords = coords[np.random.default_rng(0).permutation(len(coords))]
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    nbr, ld, _ = _setup(coords, coords, offs)
    x = torch.randn(len(coords), cin, dtype=torch.float32)
    w = torch.randn(27, cin, cout, dtype=torch.float32) / 20
    y = abi.conv_fwd(x.cuda(), w.cuda(), nbr, ld, len(coords)).cpu()
    dense = torch.zeros(1, cin, S, S, S)
    dense[0, :, coords[:, 1], coords[:, 2], coords[:, 3]] = x.T
    wt = w.reshape(3, 3, 3, cin, cout).permute(4, 3, 2, 1, 0).contiguous()
    ref = F.conv3d(dense, wt, padding=1)[0][:, coords[:, 1], coords[:, 2], coords[:, 3]].T
    assert rel_err(y.numpy(), ref.numpy()) <= 1e-5


def test_conv_linearity_and_zero_rows_large():
    """Size-independent properties at a larger size (oracle too slow): linearity
    in x and in W, and rows with no neighbours give exactly zero."""
    c = O.sphere_surface(120, seed=4)          # ~196k voxels
    far = np.array([[0, 30000, 30000, 30000], [1, -30000, 5, 5]], np.int32)   # isolated voxels
    c = np.concatenate([c, far]).astype(np.int32)
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    tin = abi.insert(c)
    nbr, ld, counts = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offs)
    n = len(c)
    g = torch.Generator().manual_seed(2)
    x1 = torch.randn(n, 64, generator=g).cuda()
    x2 = torch.randn(n, 64, generator=g).cuda()
    w = (torch.randn(27, 64, 64, generator=g) / 40).cuda()
    y1 = abi.conv_fwd(x1, w, nbr, ld, n)
    y2 = abi.conv_fwd(x2, w, nbr, ld, n)
    y12 = abi.conv_fwd(x1 + 2 * x2, w, nbr, ld, n)
    assert rel_err((y1 + 2 * y2).cpu().numpy(), y12.cpu().numpy()) <= 1e-5
    # isolated rows only see the centre offset
    yc = x1[-2:] @ w[13]
    assert rel_err(y1[-2:].cpu().numpy(), yc.cpu().numpy()) <= 1e-5
    # <dy, conv(x)> == <dgrad(dy), x> == <wgrad(x, dy), W>   (adjointness)
    dy = torch.randn(n, 64, generator=g).cuda()
    gi = abi.conv_fwd(dy, w, nbr, ld, n, transpose_w=1, kflip=1)
    gw = abi.conv_wgrad(x1, dy, nbr, ld, n, 27)
    a = float((dy.double() * y1.double()).sum())
    b = float((gi.double() * x1.double()).sum())
    cc = float((gw.double() * w.double()).sum())
    # the three inner products cancel heavily (|a| ~ sqrt(N)); compare on the scale of the summands
    scale = float(dy.double().norm() * y1.double().norm())
    assert abs(a - b) <= 1e-5 * scale and abs(a - cc) <= 1e-5 * scale


def _northstar_cloud(R, order):
    """Sphere-surface cloud in one of the row orders the engine must not care about."""
    c = O.sphere_surface(R, seed=0)
    if order == "shuffled":
        c = c[np.random.default_rng(3).permutation(len(c))]
    elif order == "blocks":      # rows grouped by 32^3 block: tiles whose rows share few offsets
        key = ((c[:, 1] >> 5).astype(np.int64) << 40) | ((c[:, 2] >> 5).astype(np.int64) << 20) | (c[:, 3] >> 5)
        c = c[np.argsort(key, kind="stable")]
    return np.ascontiguousarray(c.astype(np.int32))


@pytest.mark.parametrize("R,order", [(271, "sorted"), (120, "shuffled"), (120, "blocks")])
def test_tensor_core_path_at_scale(R, order):
    """The tcgen05 kernels at the north-star size (1 M voxels, 3^3, 64 -> 64, bf16), where the
    oracle is too slow: (1) agreement with the engine's own fp32 CUDA-core path on the same
    bf16-rounded inputs, (2) bit-identical results from run to run (no atomics, fixed reduction
    order), (3) adjointness <dy, conv(x)> = <dgrad(dy), x> = <wgrad(x, dy), W>."""
    c = _northstar_cloud(R, order)
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    tin = abi.insert(c)
    nbr, ld, _ = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offs)
    n = len(c)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(n, 64, generator=g).bfloat16().cuda()
    dy = torch.randn(n, 64, generator=g).bfloat16().cuda()
    w = (torch.randn(27, 64, 64, generator=g) / 40).bfloat16().cuda()

    def run():
        y = abi.conv_fwd(x, w, nbr, ld, n)
        assert abi.last_path() == 1, "tensor-core path expected"
        gi = abi.conv_fwd(dy, w, nbr, ld, n, transpose_w=1, kflip=1)
        gw = abi.conv_wgrad(x, dy, nbr, ld, n, 27)
        assert abi.last_path() == 1
        return y, gi, gw

    y, gi, gw = run()
    y2, gi2, gw2 = run()
    assert torch.equal(y, y2) and torch.equal(gi, gi2) and torch.equal(gw, gw2)
    # fp32 CUDA-core path on the same values
    abi.force_simt(1)
    try:
        xf, dyf, wf = x.float(), dy.float(), w.float()
        y_s = abi.conv_fwd(xf, wf, nbr, ld, n)
        gi_s = abi.conv_fwd(dyf, wf, nbr, ld, n, transpose_w=1, kflip=1)
        gw_s = abi.conv_wgrad(xf, dyf, nbr, ld, n, 27)
        assert abi.last_path() == 0
    finally:
        abi.force_simt(0)
    for got, ref in ((y, y_s), (gi, gi_s), (gw, gw_s)):
        err = float((got.float() - ref).abs().max() / ref.abs().max())
        assert err <= 1e-2, err
    a = float((dy.double() * y_s.double()).sum())
    b = float((gi_s.double() * x.double()).sum())
    cc = float((gw_s.double() * w.double()).sum())
    scale = float(dy.double().norm() * y_s.double().norm())
    assert abs(a - b) <= 1e-5 * scale and abs(a - cc) <= 1e-5 * scale


@pytest.mark.parametrize("cin,cout", [(64, 256), (256, 64), (320, 320), (64, 320), (640, 320), (320, 960), (16, 16)])
def test_tensor_core_wide_channels(cin, cout):
    """Channel tuples of the diffusion UNet (320/640/960) and the other plan shapes
    of the tcgen05 kernels (several 64-channel chunks, output split over CTAs, one tile per
    group) on 50 k voxels, with and without the tile-ordered table: agreement with the fp32
    CUDA-core path within the bf16 tolerance and bit-identical results from run to run."""
    c = _northstar_cloud(60, "shuffled")
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    tin = abi.insert(c)
    nbr, ld, _ = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offs)
    n = len(c)
    assert abi.tc_eligible(torch.bfloat16, n, 27, cin, cout)
    perm, nbr_t = abi.tile_order(nbr, n, ld, 27, tin["coords_dev"])
    g = torch.Generator().manual_seed(11)
    x = torch.randn(n, cin, generator=g).bfloat16().cuda()
    dy = torch.randn(n, cout, generator=g).bfloat16().cuda()
    w = (torch.randn(27, cin, cout, generator=g) / (27 * cin) ** 0.5).bfloat16().cuda()

    def run(ordered):
        if ordered:
            y = abi.conv_fwd_rows(x, w, nbr_t, ld, n, perm)
            gi = abi.conv_fwd_rows(dy, w, nbr_t, ld, n, perm, transpose_w=1, kflip=1)
            gw = abi.conv_wgrad(x, dy, nbr_t, ld, n, 27, dy_rows=perm)
        else:
            y = abi.conv_fwd(x, w, nbr, ld, n)
            gi = abi.conv_fwd(dy, w, nbr, ld, n, transpose_w=1, kflip=1)
            gw = abi.conv_wgrad(x, dy, nbr, ld, n, 27)
        assert abi.last_path() == 1, "tensor-core path expected"
        return y, gi, gw

    abi.force_simt(1)
    try:
        xf, dyf, wf = x.float(), dy.float(), w.float()
        ref = (abi.conv_fwd(xf, wf, nbr, ld, n), abi.conv_fwd(dyf, wf, nbr, ld, n, transpose_w=1, kflip=1),
               abi.conv_wgrad(xf, dyf, nbr, ld, n, 27))
    finally:
        abi.force_simt(0)
    for ordered in (False, True):
        a, b = run(ordered), run(ordered)
        for u, v, r in zip(a, b, ref):
            assert torch.equal(u, v)
            err = float((u.float() - r).abs().max() / r.abs().max())
            assert err <= 1e-2, (ordered, err)


@pytest.mark.parametrize("cin,cout", [(32, 64), (64, 128)])
def test_tensor_core_5x5x5(cin, cout):
    """5^3 kernel (K = 125: four mask words per tile, eight interleaved pair groups in the weight
    gradient, a 45-bit tile-order key) on the tensor-core path, plain and tile-ordered table,
    against the fp32 CUDA-core path."""
    c = _northstar_cloud(40, "shuffled")
    offs = O.kernel_offsets([5, 5, 5], [1, 1, 1])
    K = len(offs)
    tin = abi.insert(c)
    nbr, ld, _ = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offs)
    n = len(c)
    assert abi.tc_eligible(torch.bfloat16, n, K, cin, cout)
    perm, nbr_t = abi.tile_order(nbr, n, ld, K, tin["coords_dev"])
    assert np.array_equal(np.sort(perm.cpu().numpy()), np.arange(n))
    g = torch.Generator().manual_seed(17)
    x = torch.randn(n, cin, generator=g).bfloat16().cuda()
    dy = torch.randn(n, cout, generator=g).bfloat16().cuda()
    w = (torch.randn(K, cin, cout, generator=g) / (K * cin) ** 0.5).bfloat16().cuda()
    abi.force_simt(1)
    try:
        xf, dyf, wf = x.float(), dy.float(), w.float()
        ref = (abi.conv_fwd(xf, wf, nbr, ld, n), abi.conv_fwd(dyf, wf, nbr, ld, n, transpose_w=1, kflip=1),
               abi.conv_wgrad(xf, dyf, nbr, ld, n, K))
    finally:
        abi.force_simt(0)
    for ordered in (False, True):
        if ordered:
            got = (abi.conv_fwd_rows(x, w, nbr_t, ld, n, perm),
                   abi.conv_fwd_rows(dy, w, nbr_t, ld, n, perm, transpose_w=1, kflip=1),
                   abi.conv_wgrad(x, dy, nbr_t, ld, n, K, dy_rows=perm))
        else:
            got = (abi.conv_fwd(x, w, nbr, ld, n), abi.conv_fwd(dy, w, nbr, ld, n, transpose_w=1, kflip=1),
                   abi.conv_wgrad(x, dy, nbr, ld, n, K))
        assert abi.last_path() == 1
        for u, r in zip(got, ref):
            err = float((u.float() - r).abs().max() / r.abs().max())
            assert err <= 1e-2, (ordered, err)


def test_tensor_core_plans_agree_under_load(monkeypatch):
    """Regression for a pipeline race of the forward/dgrad kernel: with fewer than four tiles
    per group the A ring is deeper than the neighbour-index ring, and a gather team whose tile
    had no neighbour under an offset skipped the wait on that index slot, ran ahead and read the
    slot's next round before it was loaded (wrong rows / an illegal address, about one launch in
    thirty).  Every plan (tiles per group forced through the development knob MK_TC_T) must give
    the bit-identical result of the default plan, launch after launch."""
    c = _northstar_cloud(200, "sorted")
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    tin = abi.insert(c)
    nbr, ld, _ = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offs)
    n = len(c)
    perm, nbr_t = abi.tile_order(nbr, n, ld, 27, tin["coords_dev"])
    mask = abi.tile_mask(nbr_t, n, ld, 27)
    g = torch.Generator().manual_seed(13)
    dy = torch.randn(n, 64, generator=g).bfloat16().cuda()
    w = (torch.randn(27, 64, 64, generator=g) / 40).bfloat16().cuda()
    run = lambda: abi.conv_fwd_rows(dy, w, nbr_t, ld, n, perm, transpose_w=1, kflip=1, mask=mask)
    monkeypatch.delenv("MK_TC_T", raising=False)
    ref = run()
    assert abi.last_path() == 1
    for T, reps in (("2", 60), ("1", 30), ("3", 20), ("4", 20)):
        monkeypatch.setenv("MK_TC_T", T)
        for _ in range(reps):
            assert torch.equal(run(), ref), T


@pytest.mark.parametrize("cin,cout", [(256, 192), (512, 512)])
def test_fp32_wide_layers_on_tensor_cores_keep_1e5(cin, cout):
    """fp32 layers whose accumulation chain would exceed the ~400 MMAs a TMEM accumulator tolerates (tcgen05
    adds with truncation): the split path bounds the chain - offsets split over grid.z for forward / data
    gradient, more tile splits for the weight gradient - and adds the partials in fp32.  north_star's 1e-5
    against the fp64 kernels (themselves checked against the oracle in test_conv_fwd_bwd_same_map)."""
    c = _northstar_cloud(45, "sorted")
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    tin = abi.insert(c)
    nbr, ld, _ = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offs)
    n = len(c)
    g = torch.Generator().manual_seed(41)
    x = torch.randn(n, cin, generator=g, dtype=torch.float64).cuda()
    w = (torch.randn(27, cin, cout, generator=g, dtype=torch.float64) / np.sqrt(27 * cin)).cuda()
    dy = torch.randn(n, cout, generator=g, dtype=torch.float64).cuda()
    xf, wf, dyf = x.float(), w.float(), dy.float()
    x, w, dy = xf.double(), wf.double(), dyf.double()          # the reference sees the fp32-rounded values
    assert abi.tc_eligible(torch.float32, n, 27, cin, cout)
    rel = lambda a, b: float((a.double() - b).abs().max() / b.abs().max())
    y = abi.conv_fwd(xf, wf, nbr, ld, n)
    assert abi.last_path() == 1 and rel(y, abi.conv_fwd(x, w, nbr, ld, n)) <= 1e-5
    gi = abi.conv_fwd(dyf, wf, nbr, ld, n, transpose_w=1, kflip=1)
    assert abi.last_path() == 1 and rel(gi, abi.conv_fwd(dy, w, nbr, ld, n, transpose_w=1, kflip=1)) <= 1e-5
    gw = abi.conv_wgrad(xf, dyf, nbr, ld, n, 27)
    assert abi.last_path() == 1 and rel(gw, abi.conv_wgrad(x, dy, nbr, ld, n, 27)) <= 1e-5


def test_gather_loop_forms_agree(monkeypatch):
    """The gather warps' three loop forms for Cin = 64, Cout <= 64, K <= 32 - the general loop
    (MK_TC_NOLEAN), the lean loop (default) and the lean loop with TMA tile::gather4 for the fully
    valid 4-row groups (MK_TC_LEAN=2) - move the same bytes: bit-identical outputs, forward (plain
    and tile-ordered table) and data gradient (mirrored offsets), launch after launch."""
    c = _northstar_cloud(120, "sorted")
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    tin = abi.insert(c)
    nbr, ld, _ = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offs)
    n = len(c)
    perm, nbr_t = abi.tile_order(nbr, n, ld, 27, tin["coords_dev"])
    mask = abi.tile_mask(nbr_t, n, ld, 27)
    g = torch.Generator().manual_seed(29)
    x = torch.randn(n, 64, generator=g).bfloat16().cuda()
    runs = []
    for cout in (64, 32):
        w = (torch.randn(27, 64, cout, generator=g) / 40).bfloat16().cuda()
        runs.append(lambda w=w: abi.conv_fwd_rows(x, w, nbr_t, ld, n, perm, mask=mask))
        runs.append(lambda w=w: abi.conv_fwd(x, w, nbr, ld, n))
    wt = (torch.randn(27, 64, 64, generator=g) / 40).bfloat16().cuda()
    runs.append(lambda: abi.conv_fwd_rows(x, wt, nbr_t, ld, n, perm, transpose_w=1, kflip=1, mask=mask))
    monkeypatch.delenv("MK_TC_LEAN", raising=False)
    monkeypatch.setenv("MK_TC_NOLEAN", "1")
    refs = [r() for r in runs]
    assert abi.last_path() == 1
    monkeypatch.delenv("MK_TC_NOLEAN")
    for lean in ("1", "2"):
        monkeypatch.setenv("MK_TC_LEAN", lean)
        for _ in range(10):
            for r, ref in zip(runs, refs):
                assert torch.equal(r(), ref), lean


@pytest.mark.parametrize("R,order,with_coords", [(60, "sorted", True), (60, "shuffled", True), (40, "sorted", False)])
def test_tile_ordered_table(R, order, with_coords, monkeypatch):
    """mk_kmap_tile_order: `order` is a permutation, the ordered table is the table with its
    columns permuted (bit-exact), it touches fewer (tile, offset) pairs, and the convolution on
    it (mk_conv_fwd_rows / dy gathered for the weight gradient) gives the same result: the
    forward bit-identical (a row's accumulation sequence does not depend on its tile), the
    weight gradient within the bf16 tolerance (the voxel reduction is re-associated)."""
    c = _northstar_cloud(R, order)
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    tin = abi.insert(c)
    nbr, ld, _ = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], offs)
    n = len(c)
    perm, nbr_t = abi.tile_order(nbr, n, ld, 27, tin["coords_dev"] if with_coords else None)
    ph = perm.cpu().numpy()
    assert np.array_equal(np.sort(ph), np.arange(n))
    nh, th = nbr.cpu().numpy(), nbr_t.cpu().numpy()
    assert np.array_equal(th[:, :n], nh[:, ph]) and (th[:, n:] == -1).all()
    m0 = abi.tile_mask(nbr, n, ld, 27).cpu().numpy().view(np.uint32)
    m1 = abi.tile_mask(nbr_t, n, ld, 27).cpu().numpy().view(np.uint32)
    popc = lambda m: int(np.unpackbits(m.view(np.uint8)).sum())
    assert popc(m1) <= popc(m0)
    g = torch.Generator().manual_seed(7)
    x = torch.randn(n, 64, generator=g).bfloat16().cuda()
    dy = torch.randn(n, 64, generator=g).bfloat16().cuda()
    w = (torch.randn(27, 64, 64, generator=g) / 40).bfloat16().cuda()
    y0 = abi.conv_fwd(x, w, nbr, ld, n)
    y1 = abi.conv_fwd_rows(x, w, nbr_t, ld, n, perm)
    assert abi.last_path() == 1
    assert torch.equal(y0, y1)
    g0 = abi.conv_fwd(dy, w, nbr, ld, n, transpose_w=1, kflip=1)
    g1 = abi.conv_fwd_rows(dy, w, nbr_t, ld, n, perm, transpose_w=1, kflip=1)
    assert torch.equal(g0, g1)
    w0 = abi.conv_wgrad(x, dy, nbr, ld, n, 27)
    w1 = abi.conv_wgrad(x, abi.rows_gather(dy, perm), nbr_t, ld, n, 27)
    assert float((w0.float() - w1.float()).abs().max() / w0.float().abs().max()) <= 1e-2
    # dy read through the permutation inside the kernel (TMA gather4) = the pre-gathered copy, bit for bit
    w2 = abi.conv_wgrad(x, dy, nbr_t, ld, n, 27, dy_rows=perm)
    assert abi.last_path() == 1
    assert torch.equal(w1, w2)
    # fp64 has no tensor-core path: a row permutation must be refused, not ignored
    with pytest.raises(RuntimeError):
        abi.conv_fwd_rows(x.double(), w.double(), nbr_t, ld, n, perm)
    with pytest.raises(RuntimeError):
        abi.conv_wgrad(x.double(), dy.double(), nbr_t, ld, n, 27, dy_rows=perm)
    # fp32 off the tensor cores (MK_F32_TC=0; by default only shapes the split path cannot serve): the tuned
    # FFMA kernels honour the permutations too
    xf, wf, dyf = x.float(), w.float(), dy.float()
    monkeypatch.setenv("MK_F32_TC", "0")
    assert not abi.tc_eligible(torch.float32, n, 27, 64, 64)
    s0 = abi.conv_fwd(xf, wf, nbr, ld, n)
    s1 = abi.conv_fwd_rows(xf, wf, nbr_t, ld, n, perm, mask=abi.tile_mask(nbr_t, n, ld, 27))
    assert abi.last_path() == 0 and float((s0 - s1).abs().max() / s0.abs().max()) <= 1e-6
    s2 = abi.conv_fwd_rows(dyf, wf, nbr_t, ld, n, perm, transpose_w=1, kflip=1)
    assert float((s2 - abi.conv_fwd(dyf, wf, nbr, ld, n, transpose_w=1, kflip=1)).abs().max() / s2.abs().max()) <= 1e-6
    v0 = abi.conv_wgrad(xf, dyf, nbr, ld, n, 27)
    v1 = abi.conv_wgrad(xf, dyf, nbr_t, ld, n, 27, dy_rows=perm)
    assert float((v0 - v1).abs().max() / v0.abs().max()) <= 1e-5
    monkeypatch.delenv("MK_F32_TC")
    # fp32: where the split tensor-core path serves the shape it honours the permutations too
    x32, w32, dy32 = x.float()[:, :32].contiguous(), w.float()[:, :32].contiguous(), dy.float()
    if abi.tc_eligible(torch.float32, n, 27, 32, 64):
        f0 = abi.conv_fwd(x32, w32, nbr, ld, n)
        f1 = abi.conv_fwd_rows(x32, w32, nbr_t, ld, n, perm)
        assert abi.last_path() == 1
        assert float((f0 - f1).abs().max() / f0.abs().max()) <= 1e-5
        q0 = abi.conv_wgrad(x32, dy32, nbr, ld, n, 27)
        q1 = abi.conv_wgrad(x32, dy32, nbr_t, ld, n, 27, dy_rows=perm)
        assert abi.last_path() == 1
        assert float((q0 - q1).abs().max() / q0.abs().max()) <= 1e-5
