"""ctypes wrapper of oracle/libcpu_ref.so (C + OpenMP CPU restatement).
TEST / BASELINE INFRASTRUCTURE ONLY — see cpu_ref.c header."""
import ctypes as C

import numpy as np

from . import build_cpu_ref

_lib = None


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(build_cpu_ref.build())
        l.cr_map_create.restype = C.c_void_p
        l.cr_map_create.argtypes = [C.c_int64]
        l.cr_map_destroy.argtypes = [C.c_void_p]
        l.cr_insert_and_map.restype = C.c_int64
        l.cr_insert_and_map.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
        l.cr_kernel_map.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                    C.c_void_p, C.c_void_p]
        l.cr_conv_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_int64, C.c_void_p]
        l.cr_conv_backward.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        l.cr_num_threads.restype = C.c_int
        l.cr_set_num_threads.argtypes = [C.c_int]
        _lib = l
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class CoordMap:
    def __init__(self, coords: np.ndarray):
        self.coords_in = np.ascontiguousarray(coords, dtype=np.int32)
        n, self.ncols = self.coords_in.shape
        self.h = lib().cr_map_create(n)
        self.unique_index = np.empty(n, np.int64)
        self.inverse_map = np.empty(n, np.int64)
        nu = lib().cr_insert_and_map(self.h, _p(self.coords_in), n, self.ncols, _p(self.unique_index),
                                     _p(self.inverse_map))
        self.unique_index = self.unique_index[:nu]
        self.coords = self.coords_in[self.unique_index]

    def __del__(self):
        if getattr(self, "h", None):
            lib().cr_map_destroy(self.h)
            self.h = None


class KernelMap:
    def __init__(self, in_map: CoordMap, out_coords: np.ndarray, offsets: np.ndarray):
        out_coords = np.ascontiguousarray(out_coords, dtype=np.int32)
        offs = np.ascontiguousarray(offsets, dtype=np.int32)
        self.K, self.n_out = len(offs), len(out_coords)
        self.counts = np.zeros(self.K, np.int64)
        self.in_rows = np.empty((self.K, max(self.n_out, 1)), np.int32)
        self.out_rows = np.empty((self.K, max(self.n_out, 1)), np.int32)
        lib().cr_kernel_map(in_map.h, _p(out_coords), self.n_out, out_coords.shape[1], _p(offs), self.K,
                            _p(self.counts), _p(self.in_rows), _p(self.out_rows))

    def pairs(self):
        return [(self.in_rows[k, :c].astype(np.int64), self.out_rows[k, :c].astype(np.int64))
                for k, c in enumerate(self.counts)]


def conv_forward(x: np.ndarray, w: np.ndarray, km: KernelMap) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float32)
    w = np.ascontiguousarray(w, np.float32)
    out = np.zeros((km.n_out, w.shape[2]), np.float32)
    lib().cr_conv_forward(_p(x), _p(w), km.K, w.shape[1], w.shape[2], _p(km.counts), _p(km.in_rows),
                          _p(km.out_rows), max(km.n_out, 1), _p(out))
    return out


def conv_backward(x: np.ndarray, dy: np.ndarray, w: np.ndarray, km: KernelMap):
    x = np.ascontiguousarray(x, np.float32)
    dy = np.ascontiguousarray(dy, np.float32)
    w = np.ascontiguousarray(w, np.float32)
    gi = np.zeros_like(x)
    gw = np.zeros_like(w)
    lib().cr_conv_backward(_p(x), _p(dy), _p(w), km.K, w.shape[1], w.shape[2], _p(km.counts), _p(km.in_rows),
                           _p(km.out_rows), max(km.n_out, 1), _p(gi), _p(gw))
    return gi, gw


def num_threads() -> int:
    return int(lib().cr_num_threads())


def set_num_threads(n: int) -> None:
    """Override OMP_NUM_THREADS (torchrun exports 1 for its workers)."""
    lib().cr_set_num_threads(int(n))
