"""The C-ABI library loads and exports every symbol include/mink_b200.h declares
(no compute calls: runs without a GPU)."""
import ctypes
import os
import re

from conftest import ROOT


def _declared():
    text = open(os.path.join(ROOT, "include", "mink_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mk_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib):
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in mink_b200.h but not exported"


def test_prototypes_cover_header():
    from mink_octtree_stablediffusion_b200 import _lib
    assert sorted(_lib.PROTOTYPES) == _declared()


def test_pure_host_entry_points(lib):
    assert lib.mk_version() >= 100
    assert lib.mk_hash_capacity(1000) == 4096
    assert lib.mk_hash_capacity(0) >= 1
    assert lib.mk_insert_workspace_bytes(1 << 20) > 3 * 4 * (1 << 20)
    assert lib.mk_select_workspace_bytes(10) > 0
    assert lib.mk_kmap_pairs_workspace_bytes(1000, 27) > 0
    assert lib.mk_conv_workspace_bytes(2, 1000, 27, 64, 64) >= 27 * 64 * 64 * 4
    assert isinstance(lib.mk_last_error(), bytes)


def test_bad_arguments_return_status_not_abort(lib):
    # argument validation happens before any CUDA call
    rc = lib.mk_map_lookup(None, 1000, None, 5, 4, None, None)   # capacity not a power of two
    assert rc == 1
    assert b"power of two" in lib.mk_last_error()
    rc = lib.mk_kmap_build(None, 10, 7, None, 1024, None, 27, None, 128, None, None)
    assert rc == 1 and b"ncols" in lib.mk_last_error()
    rc = lib.mk_conv_fwd(0, None, None, 0, 0, None, None, 4, 8, 8, 27, 4, 4, None, None, None)  # ld < n_out
    assert rc == 1


def test_no_cpu_backend():
    import pytest
    import MinkowskiEngineBackend._C as _C
    with pytest.raises(RuntimeError):
        _C.CoordinateMapManagerCPU()
    assert not hasattr(_C, "ConvolutionForwardCPU")
