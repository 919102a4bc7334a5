"""`ME.MinkowskiPruning` (reference: MinkowskiPruning.py:38-121): keep the rows
of a sparse tensor where a boolean mask is true (octree decoding)."""
import torch
from torch.autograd import Function
from torch.nn import Module

from MinkowskiEngineBackend._C import CoordinateMapKey

from .MinkowskiCommon import MinkowskiModuleBase, get_minkowski_function
from .MinkowskiCoordinateManager import CoordinateManager
from .MinkowskiSparseTensor import SparseTensor


class MinkowskiPruningFunction(Function):
    @staticmethod
    def forward(ctx, in_feat: torch.Tensor, mask: torch.Tensor, in_coords_key: CoordinateMapKey,
                out_coords_key: CoordinateMapKey = None, coords_manager: CoordinateManager = None):
        ctx.in_coords_key = in_coords_key
        ctx.out_coords_key = out_coords_key
        ctx.coords_manager = coords_manager
        in_feat = in_feat.contiguous()
        fw_fn = get_minkowski_function("PruningForward", in_feat)
        return fw_fn(in_feat, mask, ctx.in_coords_key, ctx.out_coords_key, ctx.coords_manager._manager)

    @staticmethod
    def backward(ctx, grad_out_feat: torch.Tensor):
        grad_out_feat = grad_out_feat.contiguous()
        bw_fn = get_minkowski_function("PruningBackward", grad_out_feat)
        grad_in_feat = bw_fn(grad_out_feat, ctx.in_coords_key, ctx.out_coords_key, ctx.coords_manager._manager)
        return grad_in_feat, None, None, None, None


class MinkowskiPruning(MinkowskiModuleBase):
    """Remove coordinates: the output holds C = coords where mask is True and
    F = copies of those feature rows, order preserved."""

    def __init__(self):
        super().__init__()

    def forward(self, input: SparseTensor, mask: torch.Tensor):
        assert isinstance(input, SparseTensor)
        assert isinstance(mask, torch.Tensor)
        assert mask.dtype == torch.bool or mask.dtype == torch.uint8, "Mask must be a boolean tensor."
        assert mask.numel() == len(input), "Mask must have one entry per row of the input."
        out_coords_key = CoordinateMapKey(input.coordinate_map_key.get_coordinate_size())
        output = MinkowskiPruningFunction.apply(input.F, mask, input.coordinate_map_key, out_coords_key,
                                                input._manager)
        return SparseTensor(output, coordinate_map_key=out_coords_key, coordinate_manager=input._manager)

    def __repr__(self):
        return self.__class__.__name__ + "()"
