"""Run in a subprocess by tests/test_reference_l3.py: the REFERENCE's own L3 Python
(MinkowskiEngine/*.py, staged unchanged under baseline/_ref/me_l3) on top of this repo's
`MinkowskiEngineBackend._C`.  argv[1] = device ("cuda" | "cpu-oracle")."""
import sys

ROOT = sys.argv[2]
sys.path[:0] = [ROOT + "/baseline/_ref/me_l3", ROOT]
import numpy as np  # noqa: E402
import torch  # noqa: E402

import MinkowskiEngine as ME  # noqa: E402  (the reference's package)
import MinkowskiEngineBackend._C as _C  # noqa: E402  (this repo's backend)
from oracle import oracle as O  # noqa: E402

assert "/baseline/_ref/me_l3/" in ME.__file__ and "/baseline/" not in _C.__file__
dev = sys.argv[1]
if dev == "cpu-oracle":
    from oracle import me_cpu_backend as B
    ctx = B.installed()
    device = "cpu"
else:
    import contextlib
    ctx = contextlib.nullcontext()
    device = "cuda"
with ctx:
    c = np.concatenate([O.sphere_surface(12, 0, 0), O.sphere_surface(9, 1, 1)])
    torch.manual_seed(0)
    x = torch.randn(len(c), 16, device=device, requires_grad=True)
    st = ME.SparseTensor(x, torch.from_numpy(c).to(device))
    conv = ME.MinkowskiConvolution(16, 32, kernel_size=3, stride=2, dimension=3).to(device)
    up = ME.MinkowskiGenerativeConvolutionTranspose(32, 8, kernel_size=2, stride=2, dimension=3).to(device)
    y = conv(st)
    z = up(y)
    keep = z.F[:, 0] > 0
    p = ME.MinkowskiPruning()(z, keep)
    g = ME.MinkowskiGlobalAvgPooling()(p)
    g.F.sum().backward()
    # oracle restatement of the same chain
    s = O.stride_coords(c, [2, 2, 2])
    km = O.kernel_map(c, s, O.kernel_offsets([3, 3, 3], [1, 1, 1]))
    y_ref = O.conv_forward(x.detach().double().cpu().numpy(), conv.kernel.detach().double().cpu().numpy(), km, len(s))
    assert np.array_equal(O.canonical_coords(y.C.cpu().numpy()), O.canonical_coords(s))
    perm = O.row_permutation(y.C.cpu().numpy(), s)
    err = np.abs(y.F.detach().double().cpu().numpy()[perm] - y_ref).max() / np.abs(y_ref).max()
    assert err <= 1e-5, err
    fine = O.generative_coords(s, O.kernel_offsets([2, 2, 2], [1, 1, 1]))
    assert np.array_equal(O.canonical_coords(z.C.cpu().numpy()), O.canonical_coords(fine))
    assert np.array_equal(p.C.cpu().numpy(), z.C.cpu().numpy()[keep.cpu().numpy()])
    assert x.grad is not None and torch.isfinite(x.grad).all() and conv.kernel.grad.abs().sum() > 0
    km_api = st.coordinate_manager.kernel_map(st.coordinate_map_key, y.coordinate_map_key, stride=2, kernel_size=3)
    assert sum(v.shape[1] for v in km_api.values()) == sum(len(a) for a, _ in km)
print("REF-L3-OK", dev, float(err))
