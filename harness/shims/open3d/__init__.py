"""Import shim: the reference's example scripts `import open3d` at module top
(examples/ae_res.py:41-44, reconstruction.py:36) but only use it for mesh
loading and visualisation, which are outside the hot path (SURVEY.md §7.2b;
no graphics stack on this box).  Any real use raises."""


class _Unavailable:
    def __init__(self, name):
        self._name = name

    def __getattr__(self, item):
        return _Unavailable(f"{self._name}.{item}")

    def __call__(self, *args, **kwargs):
        raise RuntimeError(f"{self._name} is unavailable on this box (open3d is not installed; visualisation and "
                           "mesh IO are outside the engine's scope)")


geometry = _Unavailable("open3d.geometry")
utility = _Unavailable("open3d.utility")
io = _Unavailable("open3d.io")
visualization = _Unavailable("open3d.visualization")
__version__ = "0.17.0"  # examples/reconstruction.py:156 asserts a version >= 0.8
