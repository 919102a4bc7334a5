"""The BASELINE configurations themselves against the oracle and the frozen golden values
(tests/golden/configs_golden.npz, made by tests/golden/make_golden_configs.py):

  cfg1        batch 2 x sphere R=61, 3^3, 32 -> 64, fp32, through ME.MinkowskiConvolution (<= 1e-5)
  north_star  sphere R=271 (1.0 M voxels), 3^3, 64 -> 64, bf16: plain AND tile-ordered table,
              fwd + dgrad + wgrad vs oracle/cpu_ref.c run in fp32 on the bf16-rounded inputs
              (<= 1e-2), kernel map bit-exact (canonical pair lists, SHA-256)

CPU tests: the NumPy oracle and the C restatement reproduce the golden values.
GPU tests: the engine against the restatement run here and against the golden values."""
import os

import numpy as np
import pytest
import torch

from conftest import ROOT
from golden import configs
from oracle import cpu_ref as CR
from oracle import oracle as O

GOLD = np.load(os.path.join(ROOT, "tests", "golden", "configs_golden.npz"))
OFFS = O.kernel_offsets([3, 3, 3], [1, 1, 1])


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def _digest(name):
    return bytes(GOLD[f"{name}_kmap_sha256"]).hex()


def test_cfg1_oracles_reproduce_golden():
    c, x, w, dy = configs.inputs("cfg1")
    assert len(c) == int(GOLD["cfg1_n"])
    CR.set_num_threads(len(os.sched_getaffinity(0)))
    km = CR.KernelMap(CR.CoordMap(c), c, OFFS)
    assert configs.kmap_digest(km.pairs()) == _digest("cfg1")
    assert np.array_equal(km.counts, GOLD["cfg1_counts"])
    kmo = O.kernel_map(c, c, OFFS)
    assert configs.kmap_digest(kmo) == _digest("cfg1")
    rows = configs.sample_rows(len(c))
    y = CR.conv_forward(x, w, km)
    assert rel(y[rows], GOLD["cfg1_y_rows"]) <= 1e-6
    assert rel(y[rows], GOLD["cfg1_y_rows_oracle64"]) <= 1e-6
    gi, gw = CR.conv_backward(x, dy, w, km)
    assert rel(gi[rows], GOLD["cfg1_gi_rows"]) <= 1e-6 and rel(gw, GOLD["cfg1_gw"]) <= 1e-5


@pytest.mark.gpu
def test_cfg1_through_me_api_fp32():
    import MinkowskiEngine as ME
    c, x, w, dy = configs.inputs("cfg1")
    conv = ME.MinkowskiConvolution(32, 64, kernel_size=3, dimension=3).cuda()
    with torch.no_grad():
        conv.kernel.copy_(torch.from_numpy(w))
    xt = torch.from_numpy(x).cuda().requires_grad_(True)
    st = ME.SparseTensor(xt, torch.from_numpy(c).cuda())
    assert torch.equal(st.C.cpu(), torch.from_numpy(c))            # unique input: row order kept
    out = conv(st)
    out.F.backward(torch.from_numpy(dy).cuda())
    rows = configs.sample_rows(len(c))
    assert rel(out.F.detach().cpu().numpy()[rows], GOLD["cfg1_y_rows_oracle64"]) <= 1e-5
    assert rel(xt.grad.cpu().numpy()[rows], GOLD["cfg1_gi_rows"]) <= 1e-5
    assert rel(conv.kernel.grad.cpu().numpy(), GOLD["cfg1_gw"]) <= 1e-5
    # the whole output, not only the sampled rows, against the NumPy oracle
    km = O.kernel_map(c, c, OFFS)
    assert rel(out.F.detach().cpu().numpy(), O.conv_forward(x.astype(np.float64), w.astype(np.float64), km, len(c))) <= 1e-5
    # kernel map exported by the manager == the golden canonical map
    got = st.coordinate_manager.kernel_map(st.coordinate_map_key, st.coordinate_map_key, kernel_size=3)
    pairs = [(got[k][0].cpu().numpy(), got[k][1].cpu().numpy()) if k in got else (np.zeros(0), np.zeros(0))
             for k in range(27)]
    assert configs.kmap_digest(pairs) == _digest("cfg1")


@pytest.mark.gpu
def test_north_star_config_vs_cpu_ref_and_golden():
    import abi
    c, x, w, dy = configs.inputs("north_star")
    n = len(c)
    assert n == int(GOLD["north_star_n"]) and 990_000 < n < 1_010_000
    # the restatement, run here on the box's host cores (fp32 on the bf16-rounded values)
    CR.set_num_threads(len(os.sched_getaffinity(0)))
    km = CR.KernelMap(CR.CoordMap(c), c, OFFS)
    assert configs.kmap_digest(km.pairs()) == _digest("north_star")
    y_ref = CR.conv_forward(x, w, km)
    gi_ref, gw_ref = CR.conv_backward(x, dy, w, km)
    rows = configs.sample_rows(n)
    assert rel(y_ref[rows], GOLD["north_star_y_rows"]) <= 1e-5 and rel(gw_ref, GOLD["north_star_gw"]) <= 1e-4
    # engine: kernel map bit-exact
    tin = abi.insert(c)
    assert tin["n_unique"] == n and np.array_equal(tin["coords"], c)
    nbr, ld, counts = abi.kmap_build(tin["coords_dev"], tin["table"], tin["cap"], OFFS)
    assert np.array_equal(counts, GOLD["north_star_counts"])
    assert configs.kmap_digest(abi.kmap_pairs(nbr, n, ld, 27)) == _digest("north_star")
    xd, wd, dyd = (torch.from_numpy(t).bfloat16().cuda() for t in (x, w, dy))
    mask = abi.tile_mask(nbr, n, ld, 27)
    perm, nbr_o = abi.tile_order(nbr, n, ld, 27, tin["coords_dev"])
    mask_o = abi.tile_mask(nbr_o, n, ld, 27)
    for tag, run in (
        ("plain", lambda: (abi.conv_fwd(xd, wd, nbr, ld, n, mask=mask),
                           abi.conv_fwd(dyd, wd, nbr, ld, n, transpose_w=1, kflip=1, mask=mask),
                           abi.conv_wgrad(xd, dyd, nbr, ld, n, 27, mask=mask, dw_f32=True))),
        ("tile-ordered", lambda: (abi.conv_fwd_rows(xd, wd, nbr_o, ld, n, perm, mask=mask_o),
                                  abi.conv_fwd_rows(dyd, wd, nbr_o, ld, n, perm, transpose_w=1, kflip=1, mask=mask_o),
                                  abi.conv_wgrad(xd, dyd, nbr_o, ld, n, 27, mask=mask_o, dy_rows=perm, dw_f32=True))),
    ):
        y, gi, gw = run()
        assert abi.last_path() == 1, tag
        assert rel(y.float().cpu().numpy(), y_ref) <= 1e-2, tag
        assert rel(gi.float().cpu().numpy(), gi_ref) <= 1e-2, tag
        assert rel(gw.float().cpu().numpy(), gw_ref) <= 1e-2, tag
        assert rel(y.float().cpu().numpy()[rows], GOLD["north_star_y_rows"]) <= 1e-2, tag
        assert rel(gi.float().cpu().numpy()[rows], GOLD["north_star_gi_rows"]) <= 1e-2, tag
        assert rel(gw.float().cpu().numpy(), GOLD["north_star_gw"]) <= 1e-2, tag
