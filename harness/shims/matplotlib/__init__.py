"""Import shim (examples/ae_res.py:35 does `from matplotlib import pyplot as plt`
for validation PNG dumps only — unavailable on this box, outside the hot path)."""
__version__ = "0.0-shim"


def use(*args, **kwargs):
    pass
