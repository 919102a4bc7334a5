"""Stage the reference's WORKLOAD code next to the repo, unchanged.

`examples/*.py` (the octree VAE, the diffusion UNet, MinkUNet) and
`MinkowskiEngine/modules/*.py` (their building blocks) only use the public ME
API; north_star wants them to run unchanged on this engine.  They are copied
byte for byte into `baseline/_ref/` (git-ignored: reference sources never
enter the history; the directory still travels to the GPU box with the gpurun
snapshot).  The reference's own L3 Python (`MinkowskiEngine/*.py`,
`MinkowskiEngine/utils/`) is staged under `baseline/_ref/me_l3/` for the test
that runs it over this repo's `MinkowskiEngineBackend._C`.

`/root/reference` exists only in the build container; on the GPU box the staged
copy is what is used.
"""
from __future__ import annotations

import os
import shutil
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
REF = Path(os.environ.get("MK_REFERENCE_ROOT", "/root/reference"))
DEST = ROOT / "baseline" / "_ref"


def _copy_py(src: Path, dst: Path) -> int:
    dst.mkdir(parents=True, exist_ok=True)
    n = 0
    for f in sorted(src.glob("*.py")):
        shutil.copyfile(f, dst / f.name)
        n += 1
    return n


def is_staged() -> bool:
    return (DEST / "examples" / "ae_res.py").exists() and (DEST / "MinkowskiEngine" / "modules" / "vae_block.py").exists()


def stage(verbose: bool = False) -> bool:
    """Copy the workload files; returns True when `baseline/_ref` is usable."""
    if not (REF / "examples").is_dir():
        return is_staged()
    n = _copy_py(REF / "examples", DEST / "examples")
    n += _copy_py(REF / "MinkowskiEngine" / "modules", DEST / "MinkowskiEngine" / "modules")
    n += _copy_py(REF / "MinkowskiEngine", DEST / "me_l3" / "MinkowskiEngine")
    n += _copy_py(REF / "MinkowskiEngine" / "utils", DEST / "me_l3" / "MinkowskiEngine" / "utils")
    n += _copy_py(REF / "MinkowskiEngine" / "modules", DEST / "me_l3" / "MinkowskiEngine" / "modules")
    if verbose:
        print(f"[harness] staged {n} reference workload files into {DEST}")
    return is_staged()


if __name__ == "__main__":
    ok = stage(verbose=True)
    raise SystemExit(0 if ok else 1)
