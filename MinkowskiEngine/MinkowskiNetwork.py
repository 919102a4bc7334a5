"""`ME.MinkowskiNetwork` base class (reference: MinkowskiNetwork.py:30-57)."""
from abc import ABC, abstractmethod

import torch.nn as nn

from .MinkowskiSparseTensor import SparseTensor


class MinkowskiNetwork(nn.Module, ABC):
    """All-purpose base: stores the spatial dimension `D`."""

    def __init__(self, D):
        super().__init__()
        self.D = D

    @abstractmethod
    def forward(self, x):
        pass

    def init(self, x):
        """Initialise a SparseTensor input."""
        nrows = self.get_nrows(1)
        if nrows < 0:
            if isinstance(x, SparseTensor):
                self.initialize_coords(x.coords_man)
            else:
                raise ValueError("Initialize input coordinates")
        elif nrows != x.F.size(0):
            raise ValueError("Input size does not match the coordinate size")
