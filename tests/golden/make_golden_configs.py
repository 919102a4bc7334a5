"""Freeze golden values of the two BASELINE configurations that are single layers, so that
the NumPy oracle, the C restatement (oracle/cpu_ref.c) and the CUDA engine stay pinned to the
same bytes (run once in the build container: python tests/golden/make_golden_configs.py).

  cfg1        batch 2 x sphere R=61 (~100 k voxels), 3^3, 32 -> 64, fp32     (BASELINE configs[0])
  north_star  sphere R=271 (1.0 M voxels), 3^3, 64 -> 64, bf16-rounded inputs (BASELINE metric)

Stored per configuration: n, P, per-offset pair counts, the SHA-256 of the canonical kernel map
(per offset: in rows then out rows, ascending in the out row, int32), 64 sampled rows of y and of
grad_in, and grad_w in full (float64 from the float32 C restatement; cfg1 also from the NumPy
oracle, which must agree with it to 1e-6).  Inputs are regenerated from seeds (configs.py)."""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from golden import configs  # noqa: E402
from oracle import cpu_ref as CR  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    out = {}
    for name in ("cfg1", "north_star"):
        c, x, w, dy = configs.inputs(name)
        offs = O.kernel_offsets([3, 3, 3], [1, 1, 1])
        km = CR.KernelMap(CR.CoordMap(c), c, offs)
        pairs = km.pairs()
        y = CR.conv_forward(x, w, km)
        gi, gw = CR.conv_backward(x, dy, w, km)
        rows = configs.sample_rows(len(c))
        out[f"{name}_n"] = len(c)
        out[f"{name}_counts"] = km.counts.copy()
        out[f"{name}_kmap_sha256"] = np.frombuffer(bytes.fromhex(configs.kmap_digest(pairs)), np.uint8)
        out[f"{name}_y_rows"] = y[rows].astype(np.float64)
        out[f"{name}_gi_rows"] = gi[rows].astype(np.float64)
        out[f"{name}_gw"] = gw.astype(np.float64)
        if name == "cfg1":      # the NumPy oracle on the same inputs (float64)
            kmo = O.kernel_map(c, c, offs)
            assert configs.kmap_digest(kmo) == configs.kmap_digest(pairs)
            yo = O.conv_forward(x.astype(np.float64), w.astype(np.float64), kmo, len(c))
            assert np.abs(yo - y).max() / np.abs(yo).max() < 1e-6
            out["cfg1_y_rows_oracle64"] = yo[rows]
        print(name, len(c), int(km.counts.sum()), configs.kmap_digest(pairs)[:16])
    np.savez_compressed(os.path.join(HERE, "configs_golden.npz"), **out)


if __name__ == "__main__":
    main()
