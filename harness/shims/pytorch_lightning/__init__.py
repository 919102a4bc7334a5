"""Import shim for `pytorch_lightning` (absent from this image).  The example
scripts subclass `LightningModule` around their `nn.Module` models and call
`self.log(...)` inside `training_step`; the harness drives `training_step`,
`configure_optimizers` and the backward/optimizer step itself, which is all the
Trainer does on the hot path.  `Trainer` itself is not provided."""
import random

import numpy as np
import torch

from .core import LightningModule  # noqa: F401
from . import callbacks  # noqa: F401

__version__ = "0.0-shim"


def seed_everything(seed=None, workers=False):
    seed = 0 if seed is None else int(seed)
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    return seed


class Trainer:
    def __init__(self, *args, **kwargs):
        raise RuntimeError("pytorch_lightning.Trainer is not available (shim): drive training_step from the harness")
