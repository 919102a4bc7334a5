"""The CPU oracle against the properties the reference tree pins
and ME-independent cross-checks (dense conv3d, float64 finite differences)."""
import numpy as np
import torch
import torch.nn.functional as F

from conftest import random_cloud
from oracle import oracle as O


def test_insert_and_map_docstring_example():
    # MinkowskiCoordinateManager.py:168-176
    c = np.array([[0, 0], [0, 0], [0, 1], [0, 2]], np.int32)
    uc, ui, inv = O.insert_and_map(c)
    assert np.array_equal(c[ui], uc)
    assert np.array_equal(uc[inv], c)
    assert ui.tolist() == [0, 2, 3] and inv.tolist() == [0, 0, 1, 2]


def test_insert_and_map_first_occurrence_ascending():
    c = random_cloud(5000, 6, 3, seed=1, dup_frac=0.5)
    uc, ui, inv = O.insert_and_map(c)
    assert np.all(np.diff(ui) > 0)
    assert np.array_equal(uc[inv], c) and np.array_equal(c[ui], uc)
    assert len(np.unique(c, axis=0)) == len(uc)
    # first occurrence: no earlier row holds the same coordinate
    first = {}
    for i, r in enumerate(map(tuple, c)):
        first.setdefault(r, i)
    assert sorted(first.values()) == ui.tolist()


def test_stride_floor_division_negative():
    c = np.array([[0, -1, -2, 3], [0, -3, 1, 0], [0, 0, 0, 0], [1, -1, -1, -1]], np.int32)
    s = O.stride_coords(c, [2, 2, 2])
    assert sorted(map(tuple, s)) == sorted([(0, -2, -2, 2), (0, -4, 0, 0), (0, 0, 0, 0), (1, -2, -2, -2)])


def test_tensor_stride_multiplies():
    # docs/migration_05.md:47-50: two stride-2 on stride 3 -> 12
    c = random_cloud(300, 40, 1, seed=2) * np.array([1, 3, 3, 3], np.int32)
    s1 = O.stride_coords(c, [6, 6, 6])
    s2 = O.stride_coords(s1, [12, 12, 12])
    assert np.all(s2[:, 1:] % 12 == 0)
    assert np.array_equal(O.canonical_coords(s2), O.canonical_coords(O.stride_coords(c, [12, 12, 12])))


def test_kernel_offsets_order_and_volume():
    o = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    assert o.shape == (27, 3) and O.kernel_volume([3, 3, 3]) == 27
    assert o[0].tolist() == [-1, -1, -1] and o[1].tolist() == [0, -1, -1] and o[13].tolist() == [0, 0, 0]
    e = O.kernel_offsets([2, 2, 2], [4, 4, 4])      # even: [0, K)
    assert e.min() == 0 and e.max() == 4 and e[1].tolist() == [4, 0, 0]


def test_kernel_map_k1_is_membership():
    # examples/ae_res.py:556-570
    a = random_cloud(400, 5, 2, seed=3)
    a = np.unique(a, axis=0)
    b = np.unique(random_cloud(400, 5, 2, seed=4), axis=0)
    km = O.kernel_map(b, a, O.kernel_offsets([1, 1, 1], [1, 1, 1]))
    in_rows, out_rows = km[0]
    sb = set(map(tuple, b))
    assert sorted(out_rows.tolist()) == [i for i, r in enumerate(map(tuple, a)) if r in sb]
    assert np.array_equal(b[in_rows], a[out_rows])


def test_morton_bit_order_pinned_by_notebook():
    # .ipynb_checkpoints/Untitled1-checkpoint.ipynb cell 0: morton(32,16,0) == 139264
    code = 0
    for i in range(21):
        code |= ((32 >> i) & 1) << (3 * i + 2) | ((16 >> i) & 1) << (3 * i + 1) | ((0 >> i) & 1) << (3 * i)
    assert code == 139264
    c = np.array([[0, 32, 16, 0], [0, 0, 0, 0], [0, 1, 0, 0], [0, 0, 0, 1]], np.int32)
    assert O.morton_order(c).tolist() == [1, 3, 2, 0]


def _dense_case(S=6, cin=3, cout=5, seed=0):
    rng = np.random.default_rng(seed)
    g = np.stack(np.meshgrid(np.arange(S), np.arange(S), np.arange(S), indexing="ij"), -1).reshape(-1, 3)
    coords = np.concatenate([np.zeros((len(g), 1), np.int64), g], 1).astype(np.int32)
    coords = coords[rng.permutation(len(coords))]
    x = rng.normal(size=(len(coords), cin))
    return coords, x, rng


def test_conv_dense_grid_equals_conv3d():
    # docs/sparse_tensor_network.rst:84-89: dense special case
    S, cin, cout = 6, 3, 5
    coords, x, rng = _dense_case(S, cin, cout)
    w = rng.normal(size=(27, cin, cout))
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    km = O.kernel_map(coords, coords, offs)
    y = O.conv_forward(x, w, km, len(coords))
    dense = np.zeros((1, cin, S, S, S))
    dense[0, :, coords[:, 1], coords[:, 2], coords[:, 3]] = x
    # k = ix + 3*(iy + 3*iz): x fastest -> W[k] -> weight[co, ci, kx, ky, kz] (cross-correlation)
    wt = w.reshape(3, 3, 3, cin, cout).transpose(4, 3, 2, 1, 0)   # [co, ci, kx, ky, kz] from [kz,ky,kx,ci,co]
    ref = F.conv3d(torch.from_numpy(dense), torch.from_numpy(np.ascontiguousarray(wt)), padding=1)[0].numpy()
    assert np.allclose(y, ref[:, coords[:, 1], coords[:, 2], coords[:, 3]].T, atol=1e-10)


def test_conv_transpose_k2s2_dense_equals_conv_transpose3d():
    S, cin, cout = 4, 3, 2
    coords, x, rng = _dense_case(S, cin, cout, seed=1)
    coords = coords * np.array([1, 2, 2, 2], np.int32)             # tensor stride 2
    w = rng.normal(size=(8, cin, cout))
    offs = O.kernel_offsets([2, 2, 2], [1, 1, 1])
    fine = O.generative_coords(coords, offs)
    assert len(fine) == 8 * len(coords)
    km = O.kernel_map_transpose(coords, fine, offs)
    assert sum(len(a) for a, _ in km) == len(fine)                # every fine row has exactly one pair
    y = O.conv_forward(x, w, km, len(fine))
    dense = np.zeros((1, cin, S, S, S))
    dense[0, :, coords[:, 1] // 2, coords[:, 2] // 2, coords[:, 3] // 2] = x
    wt = w.reshape(2, 2, 2, cin, cout).transpose(3, 4, 2, 1, 0)   # [ci, co, kx, ky, kz]
    ref = F.conv_transpose3d(torch.from_numpy(dense), torch.from_numpy(np.ascontiguousarray(wt)), stride=2)[0].numpy()
    assert np.allclose(y, ref[:, fine[:, 1], fine[:, 2], fine[:, 3]].T, atol=1e-10)


def test_conv_backward_matches_finite_differences():
    coords = np.unique(random_cloud(60, 3, 2, seed=5), axis=0)
    rng = np.random.default_rng(0)
    x = rng.normal(size=(len(coords), 3))
    w = rng.normal(size=(27, 3, 2))
    g = rng.normal(size=(len(coords), 2))
    km = O.kernel_map(coords, coords, O.kernel_offsets([3, 3, 3], [1, 1, 1]))
    gi, gw = O.conv_backward(x, g, w, km)
    loss = lambda xx, ww: float((O.conv_forward(xx, ww, km, len(coords)) * g).sum())
    eps = 1e-6
    for idx in [(0, 0), (5, 2), (len(coords) - 1, 1)]:
        xp = x.copy(); xp[idx] += eps
        xm = x.copy(); xm[idx] -= eps
        assert abs((loss(xp, w) - loss(xm, w)) / (2 * eps) - gi[idx]) < 1e-5
    for idx in [(0, 0, 0), (13, 2, 1), (26, 1, 0)]:
        wp = w.copy(); wp[idx] += eps
        wm = w.copy(); wm[idx] -= eps
        assert abs((loss(x, wp) - loss(x, wm)) / (2 * eps) - gw[idx]) < 1e-5


def test_prune_keeps_order():
    # MinkowskiPruning.py:93-96
    c = random_cloud(50, 9, 1, seed=6)
    f = np.arange(100.).reshape(50, 2)
    keep = np.arange(50) % 3 == 0
    pc, pf = O.prune(c, f, keep)
    assert np.array_equal(pc, c[keep]) and np.array_equal(pf, f[keep])
    g = O.prune_backward(pf, keep, 50)
    assert np.array_equal(g[keep], pf) and np.all(g[~keep] == 0)


def test_sphere_generator_statistics():
    # N ~ 13.65 R^2 and P/N ~ 11.27 for 3^3 stride 1
    c = O.sphere_surface(40, seed=0)
    assert abs(len(c) / (13.65 * 40 * 40) - 1) < 0.05
    km = O.kernel_map(c, c, O.kernel_offsets([3, 3, 3], [1, 1, 1]))
    ratio = sum(len(a) for a, _ in km) / len(c)
    assert 10.5 < ratio < 12.0
    assert len(km[13][0]) == len(c)            # centre offset hits every row
