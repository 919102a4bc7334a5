"""Pooling / broadcast / union / cat / Morton re-sort / quantisation on the engine against the same
module run on CPU tensors over the oracle backend (float64): forward values AND backward values
(reference: MinkowskiPooling.py, MinkowskiBroadcast.py, MinkowskiOps.py:141-158,
MinkowskiTensor.py:511-546, examples/ae_res.py:92-123, utils/quantization.py:136-332)."""
import numpy as np
import pytest
import torch

from oracle import me_cpu_backend as CPUB
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp(min=1e-30))


def _cloud():
    return torch.from_numpy(np.concatenate([O.sphere_surface(10, 0, 0), O.sphere_surface(8, 1, 1), O.sphere_surface(6, 2, 2)]))


def _both(fn, c=8, seed=0, dtype=torch.float64, tol=1e-10):
    """fn(ME, SparseTensor) -> SparseTensor | Tensor; compare output rows (matched by coordinate) and the
    gradient of a random linear functional of the output with respect to the input features."""
    import MinkowskiEngine as ME
    coords = _cloud()
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(len(coords), c, generator=g, dtype=torch.float64)
    res = []
    for dev in ("cpu", "cuda"):
        xi = x.detach().clone().to(dtype).to(dev).requires_grad_(True)
        with (CPUB.installed() if dev == "cpu" else torch.enable_grad()):
            out = fn(ME, ME.SparseTensor(xi, coords.to(dev)))
            F_, C_ = (out.F, out.C.cpu().numpy()) if hasattr(out, "F") else (out, None)
            if C_ is not None and len(res) == 1 and res[0][2] is not None:
                perm = torch.from_numpy(O.row_permutation(C_, res[0][2])).to(dev)      # engine rows -> oracle order
                F_ = F_[perm]
            wgt = torch.randn(F_.shape, generator=torch.Generator().manual_seed(seed + 1), dtype=torch.float64)
            (F_ * wgt.to(F_.dtype).to(dev)).sum().backward()
        res.append((F_.detach().cpu(), xi.grad.cpu(), C_))
    assert res[0][0].shape == res[1][0].shape
    assert rel(res[1][0], res[0][0]) <= tol
    assert rel(res[1][1], res[0][1]) <= tol


@pytest.mark.parametrize("name", ["MinkowskiMaxPooling", "MinkowskiAvgPooling", "MinkowskiSumPooling"])
@pytest.mark.parametrize("ks,stride", [(2, 2), (3, 1), (3, 2)])
def test_local_pooling_fwd_bwd(name, ks, stride):
    _both(lambda ME, s: getattr(ME, name)(kernel_size=ks, stride=stride, dimension=3)(s))


def test_pooling_transpose_fwd_bwd():
    def fn(ME, s):
        down = ME.MinkowskiSumPooling(kernel_size=2, stride=2, dimension=3)(s)
        return ME.MinkowskiPoolingTranspose(kernel_size=2, stride=2, dimension=3)(down)
    _both(fn)


@pytest.mark.parametrize("name", ["MinkowskiGlobalMaxPooling", "MinkowskiGlobalAvgPooling", "MinkowskiGlobalSumPooling"])
def test_global_pooling_fwd_bwd(name):
    _both(lambda ME, s: getattr(ME, name)()(s))


@pytest.mark.parametrize("name", ["MinkowskiBroadcastAddition", "MinkowskiBroadcastMultiplication"])
def test_broadcast_fwd_bwd(name):
    def fn(ME, s):
        glob = ME.MinkowskiGlobalAvgPooling()(s)
        return getattr(ME, name)()(s, glob)
    _both(fn)


def test_broadcast_concatenation_and_cat():
    def fn(ME, s):
        glob = ME.MinkowskiGlobalSumPooling()(s)
        a = ME.MinkowskiBroadcastConcatenation()(s, glob)
        return ME.cat(a, s)
    _both(fn)


def test_union_arithmetic_different_maps():
    def fn(ME, s):
        # a second tensor on a shifted copy of part of the cloud: overlapping but different coordinate sets
        cc = s.C[: len(s) // 2].clone()
        cc[:, 1] += 1
        other = ME.SparseTensor(s.F[: len(s) // 2] * 2.0, cc, coordinate_manager=s.coordinate_manager)
        return s + other
    _both(fn)


def test_morton_resort_on_same_manager():
    """examples/ae_res.py:92-123: features / coordinates re-ordered by Morton code and inserted again
    under the same tensor stride on the same manager (needs a fresh string id)."""
    def fn(ME, s):
        down = ME.MinkowskiSumPooling(kernel_size=2, stride=2, dimension=3)(s)
        order = torch.from_numpy(O.morton_order(down.C.cpu().numpy())).to(down.F.device)
        again = ME.SparseTensor(features=down.F.index_select(0, order), coordinates=down.C.index_select(0, order),
                                tensor_stride=down.tensor_stride, coordinate_manager=down.coordinate_manager)
        assert again.coordinate_map_key != down.coordinate_map_key
        assert torch.equal(again.C, down.C.index_select(0, order))
        conv = ME.MinkowskiConvolution(8, 8, kernel_size=3, dimension=3).double().to(down.F.device)
        with torch.no_grad():
            conv.kernel.copy_(torch.randn(conv.kernel.shape, generator=torch.Generator().manual_seed(3), dtype=torch.float64))
        return conv(again)
    _both(fn)


def test_sparse_quantize_on_gpu_matches_cpu():
    import MinkowskiEngine as ME
    g = torch.Generator().manual_seed(0)
    pts = torch.rand(20000, 3, generator=g) * 30
    feats = torch.randn(20000, 4, generator=g)
    labels = torch.randint(0, 5, (20000,), generator=g)
    c_cpu, f_cpu = ME.utils.sparse_quantize(pts.numpy(), feats.numpy(), quantization_size=0.5)
    c_gpu, f_gpu = ME.utils.sparse_quantize(pts.cuda(), feats.cuda(), quantization_size=0.5)
    assert c_gpu.is_cuda and f_gpu.is_cuda
    c_cpu, f_cpu = torch.as_tensor(c_cpu), torch.as_tensor(f_cpu)
    # the kept row of every occupied cell is the first point that falls into it, on both paths
    assert torch.equal(c_gpu.cpu().int(), c_cpu.int()) and torch.equal(f_gpu.cpu(), f_cpu)
    want = np.unique(np.floor(pts.numpy() / 0.5).astype(np.int32), axis=0)
    assert len(c_gpu) == len(want)
    idx_cpu, inv_cpu = ME.utils.sparse_quantize(pts.numpy(), return_index=True, return_inverse=True, quantization_size=0.5)[-2:]
    idx_gpu, inv_gpu = ME.utils.sparse_quantize(pts.cuda(), return_index=True, return_inverse=True, quantization_size=0.5)[-2:]
    assert np.array_equal(np.asarray(idx_cpu), idx_gpu.cpu().numpy()) and np.array_equal(np.asarray(inv_cpu), inv_gpu.cpu().numpy())
    # label voting (a voxel whose points disagree gets ignore_label) on the device == host path
    _, _, l_cpu = ME.utils.sparse_quantize(pts.numpy(), feats.numpy(), labels.numpy(), quantization_size=0.5)
    _, _, l_gpu = ME.utils.sparse_quantize(pts.cuda(), feats.cuda(), labels.cuda(), quantization_size=0.5)
    assert np.array_equal(np.asarray(l_cpu), l_gpu.cpu().numpy()) and (l_gpu == -100).any()
    bc = ME.utils.batched_coordinates([c_gpu, c_gpu[:100]])
    assert bc.shape == (len(c_gpu) + 100, 4) and int(bc[-1, 0]) == 1
