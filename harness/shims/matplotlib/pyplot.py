def __getattr__(name):
    def _unavailable(*args, **kwargs):
        raise RuntimeError(f"matplotlib.pyplot.{name}: plotting is unavailable on this box (shim)")
    return _unavailable
