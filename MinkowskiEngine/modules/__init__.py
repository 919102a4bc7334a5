"""`MinkowskiEngine.modules` — model building blocks.

The reference's block files (`MinkowskiEngine/modules/{vae_block,diffusion_block,
resnet_block,...}.py`) are *workload*, not engine: they only use the public ME
API.  When the reference tree has been staged next to the repo
(`baseline/_ref/MinkowskiEngine/modules`, git-ignored) its files are importable
from here unchanged, e.g. `from MinkowskiEngine.modules.vae_block import ResNet2`.
`resnet_block` (BasicBlock / Bottleneck) is provided natively.
"""
import os

_here = os.path.dirname(__file__)
_staged = os.path.join(os.path.dirname(os.path.dirname(_here)), "baseline", "_ref", "MinkowskiEngine", "modules")
if os.path.isdir(_staged):
    __path__.append(_staged)
