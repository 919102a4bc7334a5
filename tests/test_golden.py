"""Frozen vectors (tests/golden/hotpath_small.npz, made by tests/golden/make_golden.py):
the numpy oracle, the C/OpenMP port and the CUDA path (through the C-ABI) must all
reproduce the same bytes.  The vectors are oracle outputs, not upstream-ME outputs
(the reference engine cannot run here) - "parity unpinned"."""
import os

import numpy as np
import pytest

from oracle import oracle as O

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hotpath_small.npz"))


def _split(counts, a, b):
    st = np.concatenate([[0], np.cumsum(counts)])
    return [(a[st[k]:st[k + 1]].astype(np.int64), b[st[k]:st[k + 1]].astype(np.int64)) for k in range(len(counts))]


KM3 = _split(G["km3_counts"], G["km3_in"], G["km3_out"])
KMS2 = _split(G["km_s2_counts"], G["km_s2_in"], G["km_s2_out"])


def test_oracle_reproduces_golden():
    uc, ui, inv = O.insert_and_map(G["coords"])
    assert np.array_equal(uc, G["unique_coords"]) and np.array_equal(ui, G["unique_index"])
    assert np.array_equal(inv, G["inverse"])
    su, _, _ = O.insert_and_map(O.stride_coords(uc, [2, 2, 2]))
    assert np.array_equal(su, G["stride2_coords"])
    assert np.array_equal(O.generative_coords(su, O.kernel_offsets([2, 2, 2], [1, 1, 1])), G["generative_coords"])
    km = O.kernel_map(uc, uc, O.kernel_offsets([3, 3, 3], [1, 1, 1]))
    for (a, b), (ra, rb) in zip(km, KM3):
        assert np.array_equal(a, ra) and np.array_equal(b, rb)
    assert np.allclose(O.conv_forward(G["x"], G["w"], km, len(uc)), G["y"], atol=1e-13)
    gi, gw = O.conv_backward(G["x"], G["dy"], G["w"], km)
    assert np.allclose(gi, G["grad_in"], atol=1e-13) and np.allclose(gw, G["grad_w"], atol=1e-13)
    pc, pf = O.prune(uc, G["x"], G["keep"])
    assert np.array_equal(pc, G["pruned_coords"]) and np.array_equal(pf, G["pruned_feats"])


def test_c_port_reproduces_golden():
    from oracle import cpu_ref as CR
    cmap = CR.CoordMap(G["coords"])
    assert np.array_equal(cmap.coords, G["unique_coords"]) and np.array_equal(cmap.unique_index, G["unique_index"])
    assert np.array_equal(cmap.inverse_map, G["inverse"])
    km = CR.KernelMap(cmap, G["unique_coords"], O.kernel_offsets([3, 3, 3], [1, 1, 1]))
    assert np.array_equal(np.asarray(km.counts), G["km3_counts"])
    x, w, dy = (G[k].astype(np.float32) for k in ("x", "w", "dy"))
    assert np.allclose(CR.conv_forward(x, w, km), G["y"], atol=2e-5)
    gi, gw = CR.conv_backward(x, dy, w, km)
    assert np.allclose(gi, G["grad_in"], atol=2e-5) and np.allclose(gw, G["grad_w"], atol=1e-4)


@pytest.mark.gpu
def test_cuda_path_reproduces_golden():
    import torch

    import abi
    r = abi.insert(G["coords"])
    assert np.array_equal(r["coords"], G["unique_coords"]) and np.array_equal(r["unique_index"], G["unique_index"])
    assert np.array_equal(r["inverse"], G["inverse"])
    n = r["n_unique"]
    nbr, ld, counts = abi.kmap_build(r["coords_dev"], r["table"], r["cap"], O.kernel_offsets([3, 3, 3], [1, 1, 1]))
    assert np.array_equal(counts, G["km3_counts"])
    for (a, b), (ra, rb) in zip(abi.kmap_pairs(nbr, n, ld, 27), KM3):
        assert np.array_equal(a, ra) and np.array_equal(b, rb)
    # strided k2 s2 map onto the stride-2 parents (same first-occurrence order as the oracle)
    par = abi.insert(O.stride_coords(G["unique_coords"], [2, 2, 2]))
    assert np.array_equal(par["coords"], G["stride2_coords"])
    nb2, ld2, c2 = abi.kmap_build(par["coords_dev"], r["table"], r["cap"], O.kernel_offsets([2, 2, 2], [1, 1, 1]))
    assert np.array_equal(c2, G["km_s2_counts"])
    for (a, b), (ra, rb) in zip(abi.kmap_pairs(nb2, par["n_unique"], ld2, 8), KMS2):
        assert np.array_equal(a, ra) and np.array_equal(b, rb)
    for dt, tol in ((torch.float64, 1e-11), (torch.float32, 1e-5)):       # north_star: 1e-5 for fp32
        x, w, dy = (torch.from_numpy(G[k]).to(dt).cuda() for k in ("x", "w", "dy"))
        y = abi.conv_fwd(x, w, nbr, ld, n)
        assert np.allclose(y.double().cpu().numpy(), G["y"], atol=tol, rtol=tol)
        gi = abi.conv_fwd(dy, w, nbr, ld, n, transpose_w=1, kflip=1, c_out=4)
        assert np.allclose(gi.double().cpu().numpy(), G["grad_in"], atol=tol, rtol=tol)
        gw = abi.conv_wgrad(x, dy, nbr, ld, n, 27)
        assert np.allclose(gw.double().cpu().numpy(), G["grad_w"], atol=10 * tol, rtol=tol)
