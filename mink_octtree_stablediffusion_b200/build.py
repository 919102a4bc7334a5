"""Build libmink_b200.so (C-ABI, include/mink_b200.h) in-tree with nvcc for sm_100a.

No torch headers are involved: the library takes raw device pointers, so a
plain `nvcc -shared` is all it needs.  Objects are cached by source hash under
`build/` (git-ignored); the `.so` lands in `mink_octtree_stablediffusion_b200/lib/`
so that it travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
LIB_DIR = PKG / "lib"
LIB_PATH = LIB_DIR / "libmink_b200.so"
OBJ_DIR = ROOT / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
] + os.environ.get("MK_NVCC_EXTRA", "").split()


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libmink_b200.so")


def _digest(src: Path) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    h.update(src.read_bytes())
    for hdr in sorted(CSRC.glob("*.cuh")) + [ROOT / "include" / "mink_b200.h"]:
        h.update(hdr.read_bytes())
    return h.hexdigest()[:16]


def _compile(src: Path, verbose: bool) -> Path:
    obj = OBJ_DIR / f"{src.stem}.{_digest(src)}.o"
    if obj.exists():
        return obj
    for old in OBJ_DIR.glob(f"{src.stem}.*.o"):
        old.unlink()
    cmd = [_nvcc(), *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = OBJ_DIR / f"{src.stem}.ptxas.log"
    log.write_text(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src.name}:\n{res.stdout}\n{res.stderr}")
    if verbose:
        print(f"[mink_b200] compiled {src.name}", file=sys.stderr)
    return obj


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def have_nvcc() -> bool:
    try:
        _nvcc()
        return True
    except RuntimeError:
        return False


def stale_reason() -> str:
    """'' if the library's stamp names the digests of the sources (and compiler flags) as they are now.
    Development builds made with other flags (MK_NVCC_EXTRA=-DMK_TC_PROF ...) are run with
    MK_ALLOW_STALE_LIB=1."""
    stamp = LIB_DIR / "libmink_b200.stamp"
    if not stamp.exists():
        return "no stamp"
    have = stamp.read_text()
    want = ",".join(f"{s.stem}.{_digest(s)}.o" for s in sources())
    if have == want:
        return ""
    if set(n.split(".")[0] for n in have.split(",")) != set(s.stem for s in sources()):
        return "source files added or removed"
    return "" if os.environ.get("MK_ALLOW_STALE_LIB") else "sources changed since it was linked"


def build_library(force: bool = False, verbose: bool = False) -> Path:
    """Compile every csrc/*.cu for sm_100a and link the shared library."""
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    if force:
        for old in OBJ_DIR.glob("*.o"):
            old.unlink()
    srcs = sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), srcs))
    stamp = LIB_DIR / "libmink_b200.stamp"
    want = ",".join(o.name for o in objs)
    if LIB_PATH.exists() and stamp.exists() and stamp.read_text() == want and not force:
        return LIB_PATH
    cmd = [_nvcc(), "-shared", "-o", str(LIB_PATH), *map(str, objs),
           "-gencode", "arch=compute_100a,code=sm_100a"]  # cudart static; no libcuda link
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    stamp.write_text(want)
    if verbose:
        print(f"[mink_b200] linked {LIB_PATH}", file=sys.stderr)
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True))
