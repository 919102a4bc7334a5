"""B200-native sparse-tensor convolution engine (hot path of
dreaming-huang/Mink-OctTree-StableDiffusion's MinkowskiEngine fork).

Layout:
  csrc/         hand-written sm_100a CUDA kernels + the C-ABI (include/mink_b200.h)
  _lib.py       ctypes binding of libmink_b200.so
  backend*.py   host-side mirror of the reference's `MinkowskiEngineBackend._C`
  build.py      in-tree nvcc build

`MinkowskiEngineBackend._C` and `MinkowskiEngine` at the repo root re-export
this package under the reference's import names.
"""
from . import _lib  # noqa: F401

__version__ = "0.1.0"


def load_library():
    """Load libmink_b200.so (builds it in-tree with nvcc when missing)."""
    return _lib.load()
