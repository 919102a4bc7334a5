"""Synthetic workloads of SURVEY.md §8d and loaders for the reference's
UNCHANGED model code (staged by harness/stage.py).

Nothing here is engine code: it builds inputs (G3 voxelised implicit surfaces),
imports `examples.ae_res.VAE` / `examples.diffusion.UNet` / `examples.minkunet.
MinkUNet34C` from the staged reference files, and drives the reference's own
`training_step`.
"""
from __future__ import annotations

import importlib
import os
import sys
import types
from argparse import Namespace
from pathlib import Path
from typing import List

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
SHIMS = Path(__file__).resolve().parent / "shims"
STAGED = ROOT / "baseline" / "_ref"


def setup_paths() -> None:
    """repo root (this engine's `MinkowskiEngine`), then the import shims, then
    the staged reference tree (`examples` package)."""
    from . import stage
    if not stage.is_staged() and not stage.stage():
        raise RuntimeError("reference workload files are not staged: run `python harness/stage.py` in the build "
                           "container (needs /root/reference)")
    for p in (str(STAGED), str(SHIMS), str(ROOT)):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)


# ---------------------------------------------------------------------------
# G3: voxelised implicit surfaces (sphere / torus / gyroid), per-sample seed = index
# ---------------------------------------------------------------------------
def g3_shape(res: int, index: int) -> np.ndarray:
    """int32 [N, 3] voxel coordinates in [0, res)^3 of the surface of shape `index`
    (index % 3: sphere r = 0.45 res, torus, gyroid sheet), jittered by the seed."""
    rng = np.random.default_rng(index)
    kind = index % 3
    centre = res / 2 + rng.uniform(-0.02, 0.02, size=3) * res
    out: List[np.ndarray] = []
    ax = np.arange(res, dtype=np.float32) + 0.5
    slab = max(1, (1 << 22) // (res * res))
    for z0 in range(0, res, slab):
        z = ax[z0:z0 + slab]
        X, Y, Z = np.meshgrid(ax, ax, z, indexing="ij")
        x, y, zz = X - centre[0], Y - centre[1], Z - centre[2]
        if kind == 0:
            d = np.sqrt(x * x + y * y + zz * zz) - 0.45 * res
            hit = np.abs(d) < 0.55
        elif kind == 1:
            R, r = 0.32 * res, 0.12 * res
            q = np.sqrt(x * x + y * y) - R
            d = np.sqrt(q * q + zz * zz) - r
            hit = np.abs(d) < 0.55
        else:
            w = 2 * np.pi * 2.0 / res          # two periods across the grid
            g = np.sin(w * X) * np.cos(w * Y) + np.sin(w * Y) * np.cos(w * Z) + np.sin(w * Z) * np.cos(w * X)
            inside = (x * x + y * y + zz * zz) < (0.46 * res) ** 2
            hit = (np.abs(g) < 1.1 * w) & inside
        idx = np.nonzero(hit)
        out.append(np.stack([idx[0], idx[1], idx[2] + z0], axis=1).astype(np.int32))
    return np.concatenate(out, axis=0)


def g3_batch(res: int, batch: int, first_index: int = 0) -> List[np.ndarray]:
    return [g3_shape(res, first_index + i) for i in range(batch)]


def batched(coords_list: List[np.ndarray]) -> np.ndarray:
    """int32 [N, 4], batch index first (utils/collation.py:89-91)."""
    return np.concatenate([np.concatenate([np.full((len(c), 1), b, np.int32), c], axis=1)
                           for b, c in enumerate(coords_list)], axis=0)


# ---------------------------------------------------------------------------
# the reference's models, unchanged
# ---------------------------------------------------------------------------
def vae_config(resolution=128, batch_size=16, vae_channel=(32, 128, 512, 512, 4), **over) -> Namespace:
    """argparse defaults of examples/ae_res.py:955-974 (with the config-2 values)."""
    cfg = Namespace(max_epochs=1000, lr=1e-2, batch_size=batch_size, ngpus=1, num_workers=0,
                    resolution=resolution, window_size=50, floder="ae_v100", max_batch_len=2000000, recover=None,
                    seed=42, save_every=20, cache=False, noise_point_max=1000, noise_point_mode="none",
                    noise_near=False, vae_channel=tuple(vae_channel))
    for k, v in over.items():
        setattr(cfg, k, v)
    return cfg


def load_vae(cfg: Namespace, module: str = "examples.ae_res"):
    """(VAE instance, VaeModule instance, module) from the staged, unmodified file.
    The file reads its argparse namespace through a module-level global `config`
    that `__main__` sets (examples/ae_res.py:954,976): the harness sets it."""
    setup_paths()
    m = importlib.import_module(module)
    m.config = cfg
    model = m.VAE(cfg)
    pl_module = m.VaeModule(model, cfg)
    return model, pl_module, m


def diffusion_config(resolution=256, batch_size=4, unet_channel=(4, 320, 640, 960), **over) -> Namespace:
    cfg = Namespace(lr=1e-4, batch_size=batch_size, ngpus=1, num_workers=0, resolution=resolution, window_size=50,
                    max_batch_len=2000000, seed=42, noise_point_max=1000, noise_point_mode="none", noise_near=False,
                    vae_channel=(32, 128, 512, 512, 4), unet_channel=tuple(unet_channel), ddpm_num_steps=1000,
                    ddpm_beta_schedule="scaled_linear", prediction_type="epsilon", vae_scale=0.1428,
                    time_embedding_norm="default", group=1, warmup=100, with_attn=False, floder="diff")
    for k, v in over.items():
        setattr(cfg, k, v)
    return cfg


def load_unet(cfg: Namespace):
    setup_paths()
    if "transformers" not in sys.modules:
        # examples/diffusion.py:60 imports CLIP classes at module top and never uses them on this path;
        # importing the real `transformers` costs ~20 s, a stand-in with the three names is enough
        stub = types.ModuleType("transformers")
        for n in ("CLIPTextModel", "CLIPTokenizer", "CLIPVisionModel", "CLIPProcessor", "CLIPModel",
                  "CLIPImageProcessor"):
            setattr(stub, n, type(n, (), {}))
        sys.modules["transformers"] = stub
    for name in ("examples.ae_res", "examples.ae_res_v100", "examples.dataset"):
        importlib.import_module(name).config = cfg
    m = importlib.import_module("examples.diffusion")
    m.config = cfg
    return m


def load_minkunet():
    """examples/resnet.py:41-43 downloads `1.ply` at import time when it is not in the
    working directory (no network here, and the file is only used by its demo
    `__main__`): the download call is a no-op while the module is imported."""
    setup_paths()
    import urllib.request
    real = urllib.request.urlretrieve
    urllib.request.urlretrieve = lambda *a, **k: None
    try:
        return importlib.import_module("examples.minkunet")
    finally:
        urllib.request.urlretrieve = real
