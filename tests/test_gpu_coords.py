"""GPU parity (through the C-ABI) of the integer stages against the CPU oracle:
hash insert/lookup, stride, generative expansion, kernel maps.  Bit-exact."""
import numpy as np
import pytest
import torch

import abi
from conftest import random_cloud
from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,span,batch,dup,ncols", [
    (0, 4, 1, 0.0, 4), (1, 4, 1, 0.0, 4), (4, 2, 1, 0.9, 2), (1000, 3, 2, 0.5, 4), (5000, 300, 4, 0.1, 4),
    (70001, 25, 3, 0.3, 4), (3000, 40, 2, 0.2, 3), (200000, 32000, 16, 0.05, 4)])
def test_insert_matches_oracle_exactly(n, span, batch, dup, ncols):
    c = random_cloud(n, span, batch, seed=n + ncols, ncols=ncols, dup_frac=dup)
    r = abi.insert(c)
    uc, ui, inv = O.insert_and_map(c)
    assert r["status"] == 0 and r["n_unique"] == len(uc)
    # exact row order: first occurrence, ascending
    assert np.array_equal(r["unique_index"], ui)
    assert np.array_equal(r["inverse"], inv)
    assert np.array_equal(r["coords"], uc)
    # every inserted coordinate is found at its compact row; absent ones give -1
    if n:
        assert np.array_equal(abi.lookup(r["table"], r["cap"], c), inv)
        miss = c.copy()
        miss[:, 1] += 2 * span + 5
        assert np.all(abi.lookup(r["table"], r["cap"], miss) == -1)


def test_insert_docstring_example():
    r = abi.insert(np.array([[0, 0], [0, 0], [0, 1], [0, 2]], np.int32))
    assert r["unique_index"].tolist() == [0, 2, 3] and r["inverse"].tolist() == [0, 0, 1, 2]


def test_insert_range_limits():
    ok = np.array([[65534, 32767, -32768, 0], [0, -32768, 32767, 32767]], np.int32)
    r = abi.insert(ok)
    assert r["status"] == 0 and r["n_unique"] == 2
    assert abi.lookup(r["table"], r["cap"], ok).tolist() == [0, 1]
    bad = np.array([[0, 40000, 0, 0]], np.int32)
    assert abi.insert(bad)["status"] == 2          # MK_ERANGE, reported not trapped


def test_insert_many_duplicates_of_one_key():
    c = np.zeros((100000, 4), np.int32)
    c[50000:, 1] = 7
    r = abi.insert(c)
    assert r["n_unique"] == 2 and r["unique_index"].tolist() == [0, 50000]
    assert np.array_equal(r["inverse"], (np.arange(100000) >= 50000).astype(np.int32))


@pytest.mark.parametrize("stride", [[2, 2, 2], [4, 4, 4], [2, 1, 3]])
def test_stride_matches_oracle(lib, stride):
    from mink_octtree_stablediffusion_b200 import _lib
    c = np.unique(random_cloud(20000, 40, 3, seed=11), axis=0)
    src = torch.from_numpy(c).cuda()
    out = torch.empty_like(src)
    ts = np.asarray(stride, np.int32)
    _lib.check(lib.mk_coords_stride(abi.P(src), len(c), 4, abi.HP(ts), abi.P(out), abi.S()))
    raw = out.cpu().numpy()
    exp = c.astype(np.int64).copy()
    exp[:, 1:] = np.floor_divide(exp[:, 1:], ts[None]) * ts[None]
    assert np.array_equal(raw, exp.astype(np.int32))
    r = abi.insert(raw)
    assert np.array_equal(O.canonical_coords(r["coords"]), O.canonical_coords(O.stride_coords(c, stride)))
    assert np.array_equal(r["coords"], O.stride_coords(c, stride))    # same first-occurrence order


def test_expand_k2s2_octree_children(lib):
    from mink_octtree_stablediffusion_b200 import _lib
    c = np.unique(random_cloud(5000, 20, 2, seed=12), axis=0) * np.array([1, 2, 2, 2], np.int32)
    offs = O.kernel_offsets([2, 2, 2], [1, 1, 1]).astype(np.int32)
    src = torch.from_numpy(c).cuda()
    out = torch.empty((len(c) * 8, 4), dtype=torch.int32, device="cuda")
    _lib.check(lib.mk_coords_expand(abi.P(src), len(c), 4, abi.HP(offs), 8, abi.P(out), abi.S()))
    r = abi.insert(out.cpu().numpy())
    assert r["n_unique"] == 8 * len(c)             # children of distinct parents never collide
    assert np.array_equal(O.canonical_coords(r["coords"]), O.canonical_coords(O.generative_coords(c, offs)))


def _check_kmap(in_c, out_c, offs):
    tin = abi.insert(in_c)
    assert tin["n_unique"] == len(in_c)
    out_dev = torch.from_numpy(out_c).cuda()
    nbr, ld, counts = abi.kmap_build(out_dev, tin["table"], tin["cap"], offs)
    ref = O.kernel_map(in_c, out_c, offs)
    assert counts.tolist() == [len(a) for a, _ in ref]
    pairs = abi.kmap_pairs(nbr, len(out_c), ld, len(offs))
    for k, ((ir, orr), (rir, rorr)) in enumerate(zip(pairs, ref)):
        assert np.array_equal(orr, rorr), k          # ascending in the output row
        assert np.array_equal(ir, rir), k
    # dense table agrees with the pairs
    h = nbr.cpu().numpy()
    for k, (rir, rorr) in enumerate(ref):
        col = np.full(len(out_c), -1, np.int64)
        col[rorr] = rir
        assert np.array_equal(h[k, :len(out_c)], col)
    return nbr, ld, ref


@pytest.mark.parametrize("ks,scale", [([3, 3, 3], [1, 1, 1]), ([2, 2, 2], [1, 1, 1]), ([5, 5, 5], [1, 1, 1]),
                                      ([3, 3, 3], [2, 2, 2]), ([1, 1, 1], [1, 1, 1]), ([3, 1, 5], [1, 2, 1])])
def test_kernel_map_same_map(ks, scale):
    c = np.unique(random_cloud(6000, 9, 2, seed=13), axis=0) * np.array([1] + scale, np.int32)
    _check_kmap(c, c, O.kernel_offsets(ks, scale))


@pytest.mark.parametrize("ks", [[2, 2, 2], [3, 3, 3], [4, 4, 4], [5, 3, 3]])
def test_kernel_map_at_coordinate_limits(ks):
    """Rows on the int16 borders: neighbours that would leave the key range are
    absent (never aliased onto the opposite border), in both build kernels."""
    rng = np.random.default_rng(5)
    edge = np.array([-32768, -32767, -32766, -32765, -32764, 32763, 32764, 32765, 32766, 32767, -1, 0, 1], np.int64)
    c = np.stack([rng.integers(0, 2, 4000), rng.choice(edge, 4000), rng.choice(edge, 4000),
                  rng.choice(edge, 4000)], 1).astype(np.int32)
    c[:200, 0] = 65534
    c = np.unique(c, axis=0)
    offs = O.kernel_offsets(ks, [1, 1, 1])
    _check_kmap(c, c, offs)             # row-structured kernel
    _check_kmap(c, c, -offs)            # mirrored order (transposed maps): general kernel


@pytest.mark.parametrize("ncols,ks", [(2, [3]), (2, [5]), (3, [3, 3]), (3, [2, 2]), (3, [3, 1])])
def test_kernel_map_low_dimension(ncols, ks):
    c = np.unique(random_cloud(3000, 40 if ncols == 2 else 12, 3, seed=17, ncols=ncols), axis=0)
    _check_kmap(c, c, O.kernel_offsets(ks, [1] * (ncols - 1)))


def test_kernel_map_strided_and_inverse():
    c = O.sphere_surface(30, seed=3)
    s = O.stride_coords(c, [2, 2, 2])
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    nbr, ld, ref = _check_kmap(c, s, offs)
    inv, ld_in = abi.kmap_invert(nbr, len(s), ld, 27, len(c))
    hi = inv.cpu().numpy()
    for k, (rir, rorr) in enumerate(ref):
        col = np.full(len(c), -1, np.int64)
        col[rir] = rorr
        assert np.array_equal(hi[k, :len(c)], col)


def test_kernel_map_sphere_counts():
    c = O.sphere_surface(61, seed=0)
    nbr, ld, ref = _check_kmap(c, c, O.kernel_offsets([3, 3, 3], [1, 1, 1]))
    assert abs(sum(len(a) for a, _ in ref) / len(c) - 11.27) < 0.5


def test_kernel_map_empty():
    c = np.zeros((0, 4), np.int32)
    one = np.array([[0, 1, 2, 3]], np.int32)
    tin = abi.insert(one)
    nbr, ld, counts = abi.kmap_build(torch.from_numpy(c).cuda(), tin["table"], tin["cap"],
                                     O.kernel_offsets([3, 3, 3], [1, 1, 1]))
    assert counts.sum() == 0


@pytest.mark.parametrize("n,dup", [(1, 0.0), (2047, 0.3), (2049, 0.0), (300_000, 0.5), (1_000_003, 0.05), (70_000, 0.999)])
def test_insert2_matches_insert_and_oracle(n, dup):
    """mk_map_insert2 (one-pass dedup / ordered compaction, decoupled look-back) == mk_map_insert == oracle."""
    c = random_cloud(n, 40 if dup > 0.9 else 300, 3, seed=n, dup_frac=dup)
    a = abi.insert(c)
    for i64 in (False, True):
        b = abi.insert2(c, int64_maps=i64)
        assert b["status"] == 0 and b["n_unique"] == a["n_unique"]
        assert np.array_equal(b["unique_index"], a["unique_index"])
        assert np.array_equal(b["inverse"], a["inverse"])
        assert np.array_equal(b["coords"], a["coords"])
        q = c[:: max(1, n // 5000)]
        assert np.array_equal(abi.lookup(b["table"], b["cap"], q), abi.lookup(a["table"], a["cap"], q))
    u, ui, inv = O.insert_and_map(c)
    assert np.array_equal(b["coords"], u) and np.array_equal(b["unique_index"], ui) and np.array_equal(b["inverse"], inv)
    # rows known to be distinct: no dedup pass, ids = row numbers
    d = abi.insert2(u, unique=True)
    assert d["n_unique"] == len(u) and np.array_equal(d["coords"], u)
    assert np.array_equal(d["unique_index"], np.arange(len(u))) and np.array_equal(d["inverse"], np.arange(len(u)))
    assert np.array_equal(abi.lookup(d["table"], d["cap"], u), np.arange(len(u)))


def test_octree_children_map_closed_form_equals_probed_map():
    """mk_kmap_children (no probes) == the probed transposed kernel map of the same octree split, and the
    ME path that uses it gives the oracle's generative coordinates and features."""
    import MinkowskiEngine as ME
    from mink_octtree_stablediffusion_b200 import _lib
    lib = _lib.load()
    c = O.sphere_surface(40, 0, 0)
    coarse = O.stride_coords(np.concatenate([c, O.sphere_surface(25, 1, 1)]), [4, 4, 4])
    offs = O.kernel_offsets([2, 2, 2], [2, 2, 2])
    fine = (coarse[:, None, :].astype(np.int64) + np.concatenate([np.zeros((8, 1), np.int64), offs], 1)[None]).reshape(-1, 4)
    tin = abi.insert(coarse)
    nbr_p, ld, _ = abi.kmap_build(torch.from_numpy(fine.astype(np.int32)).cuda(), tin["table"], tin["cap"], -offs)
    n_in, n_out = len(coarse), len(fine)
    ld_in = max(128, (n_in + 127) // 128 * 128)
    nbr = torch.empty((8, ld), dtype=torch.int32, device="cuda")
    inv = torch.empty((8, ld_in), dtype=torch.int32, device="cuda")
    order = torch.empty(n_out, dtype=torch.int32, device="cuda")
    nbr_o = torch.empty((8, ld), dtype=torch.int32, device="cuda")
    _lib.check(lib.mk_kmap_children(n_in, 8, ld, ld_in, abi.P(nbr), abi.P(inv), abi.P(order), abi.P(nbr_o), abi.S()))
    assert torch.equal(nbr, nbr_p)
    inv_p, _ = abi.kmap_invert(nbr_p, n_out, ld, 8, n_in)
    assert torch.equal(inv, inv_p)
    o = order.long()
    assert torch.equal(torch.sort(o)[0], torch.arange(n_out, device="cuda"))
    assert torch.equal(nbr_o[:, :n_out], nbr[:, :n_out][:, o]) and bool((nbr_o[:, n_out:] == -1).all())
    # through the ME API: generative transpose k2 s2 (bf16: tile-ordered closed form; fp32: plain)
    ref_km = O.kernel_map_transpose(coarse, fine.astype(np.int32), offs)
    for dtype, tol in ((torch.float32, 1e-5), (torch.bfloat16, 1e-2)):
        torch.manual_seed(0)
        up = ME.MinkowskiGenerativeConvolutionTranspose(32, 16, kernel_size=2, stride=2, dimension=3).to(dtype).cuda()
        x = torch.randn(n_in, 32).to(dtype).cuda().requires_grad_(True)
        st = ME.SparseTensor(x, torch.from_numpy(coarse).cuda(), tensor_stride=4)
        out = up(st)
        assert out.tensor_stride == [2, 2, 2] and np.array_equal(out.C.cpu().numpy(), fine)
        dy = torch.randn(n_out, 16).to(dtype).cuda()
        out.F.backward(dy)
        xr, wr, dyr = (t.detach().double().cpu().numpy() for t in (x, up.kernel, dy))
        y_ref = O.conv_forward(xr, wr, ref_km, n_out)
        gi_ref, gw_ref = O.conv_backward(xr, dyr, wr, ref_km)
        rel = lambda a, b: float(np.abs(a.detach().double().cpu().numpy() - b).max() / np.abs(b).max())  # noqa: E731
        assert rel(out.F, y_ref) <= tol and rel(x.grad, gi_ref) <= tol and rel(up.kernel.grad, gw_ref) <= tol
