"""`ME.MinkowskiUnion` (reference: MinkowskiUnion.py:33-156): feature sum over
the union of the input coordinate maps."""
import torch
from torch.nn import Module

from MinkowskiEngineBackend._C import CoordinateMapKey

from .MinkowskiSparseTensor import SparseTensor


class MinkowskiUnion(Module):
    """A = B + C + ... on the union of coordinates (missing entries count as 0)."""

    def __init__(self):
        super().__init__()

    def forward(self, *inputs):
        assert isinstance(inputs, (list, tuple)), "The input must be a list or tuple"
        for s in inputs:
            assert isinstance(s, SparseTensor), "Inputs must be sparse tensors."
        assert len(inputs) > 1, "input must be a set with at least 2 SparseTensors"
        ref_mgr = inputs[0].coordinate_manager
        for s in inputs:
            assert ref_mgr == s.coordinate_manager, \
                "Invalid coordinate manager. All inputs must have the same coordinate manager."
        out_key = CoordinateMapKey(inputs[0].coordinate_map_key.get_coordinate_size())
        maps = ref_mgr.union_map([s.coordinate_map_key for s in inputs], out_key)
        n_out = ref_mgr.size(out_key)
        out = torch.zeros((n_out, inputs[0].F.size(1)), dtype=inputs[0].dtype, device=inputs[0].device)
        for s, m in zip(inputs, maps):
            out = out.index_add(0, m[1], s.F[m[0]])
        return SparseTensor(out, coordinate_map_key=out_key, coordinate_manager=ref_mgr)

    def __repr__(self):
        return self.__class__.__name__ + "()"
