"""The C + OpenMP CPU restatement (timed CPU baseline) against the NumPy oracle."""
import numpy as np

from conftest import random_cloud
from oracle import cpu_ref as R
from oracle import oracle as O


def test_insert_matches_numpy_oracle():
    c = random_cloud(20000, 9, 3, seed=31, dup_frac=0.3)
    m = R.CoordMap(c)
    uc, ui, inv = O.insert_and_map(c)
    assert np.array_equal(m.unique_index, ui) and np.array_equal(m.inverse_map, inv)
    assert np.array_equal(m.coords, uc)


def test_kernel_map_and_conv_match_numpy_oracle():
    c = O.sphere_surface(20, seed=7)
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    m = R.CoordMap(c)
    km = R.KernelMap(m, c, offs)
    ref = O.kernel_map(c, c, offs)
    for (a, b), (ra, rb) in zip(km.pairs(), ref):
        assert np.array_equal(a, ra) and np.array_equal(b, rb)
    rng = np.random.default_rng(0)
    x = rng.normal(size=(len(c), 32)).astype(np.float32)
    w = (rng.normal(size=(27, 32, 64)) / 30).astype(np.float32)
    dy = rng.normal(size=(len(c), 64)).astype(np.float32)
    y = R.conv_forward(x, w, km)
    y_ref = O.conv_forward(x.astype(np.float64), w.astype(np.float64), ref, len(c))
    assert np.abs(y - y_ref).max() <= 1e-5 * np.abs(y_ref).max()
    gi, gw = R.conv_backward(x, dy, w, km)
    gi_ref, gw_ref = O.conv_backward(x.astype(np.float64), dy.astype(np.float64), w.astype(np.float64), ref)
    assert np.abs(gi - gi_ref).max() <= 1e-5 * np.abs(gi_ref).max()
    assert np.abs(gw - gw_ref).max() <= 1e-4 * np.abs(gw_ref).max()


def test_strided_kernel_map():
    c = O.sphere_surface(15, seed=8)
    s = O.stride_coords(c, [2, 2, 2])
    offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
    km = R.KernelMap(R.CoordMap(c), s, offs)
    for (a, b), (ra, rb) in zip(km.pairs(), O.kernel_map(c, s, offs)):
        assert np.array_equal(a, ra) and np.array_equal(b, rb)
