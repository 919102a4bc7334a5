"""Broadcast layers (reference: MinkowskiBroadcast.py:40-253): combine a sparse
tensor with per-batch global features (rows of the origin map)."""
import torch
from torch.autograd import Function
from torch.nn import Module

from MinkowskiEngineBackend._C import BroadcastMode, CoordinateMapKey

from .MinkowskiCommon import MinkowskiModuleBase, get_minkowski_function
from .MinkowskiCoordinateManager import CoordinateManager
from .MinkowskiSparseTensor import SparseTensor


class MinkowskiBroadcastFunction(Function):
    @staticmethod
    def forward(ctx, input_features: torch.Tensor, input_features_global: torch.Tensor,
                operation_type: BroadcastMode, in_coords_key: CoordinateMapKey,
                glob_coords_key: CoordinateMapKey, coords_manager: CoordinateManager):
        assert isinstance(operation_type, BroadcastMode)
        ctx.saved_vars = (input_features, input_features_global, operation_type, in_coords_key, glob_coords_key,
                          coords_manager)
        fw_fn = get_minkowski_function("BroadcastForward", input_features)
        return fw_fn(input_features.contiguous(), input_features_global.contiguous(), operation_type,
                     in_coords_key, glob_coords_key, coords_manager._manager)

    @staticmethod
    def backward(ctx, grad_out_feat):
        grad_out_feat = grad_out_feat.contiguous()
        x, g, operation_type, in_coords_key, glob_coords_key, coords_manager = ctx.saved_vars
        bw_fn = get_minkowski_function("BroadcastBackward", grad_out_feat)
        grad_in_feat, grad_in_feat_glob = bw_fn(x.contiguous(), g.contiguous(), grad_out_feat, operation_type,
                                                in_coords_key, glob_coords_key, coords_manager._manager)
        return grad_in_feat, grad_in_feat_glob, None, None, None, None


class MinkowskiBroadcastBase(MinkowskiModuleBase):
    def __init__(self, operation_type):
        MinkowskiModuleBase.__init__(self)
        assert isinstance(operation_type, BroadcastMode)
        self.operation_type = operation_type
        self.broadcast = MinkowskiBroadcastFunction

    def forward(self, input: SparseTensor, input_glob: SparseTensor):
        assert isinstance(input, SparseTensor)
        output = self.broadcast.apply(input.F, input_glob.F, self.operation_type, input.coordinate_map_key,
                                      input_glob.coordinate_map_key, input.coordinate_manager)
        return SparseTensor(output, coordinate_map_key=input.coordinate_map_key,
                            coordinate_manager=input.coordinate_manager)

    def __repr__(self):
        return self.__class__.__name__


class MinkowskiBroadcastAddition(MinkowskiBroadcastBase):
    """y_{i, b} = x_{i, b} + g_b"""

    def __init__(self):
        MinkowskiBroadcastBase.__init__(self, BroadcastMode.ELEMENTWISE_ADDITON)


class MinkowskiBroadcastMultiplication(MinkowskiBroadcastBase):
    """y_{i, b} = x_{i, b} * g_b"""

    def __init__(self):
        MinkowskiBroadcastBase.__init__(self, BroadcastMode.ELEMENTWISE_MULTIPLICATION)


class MinkowskiBroadcast(Module):
    """Expand the per-batch global features to the rows of `input`."""

    def __repr__(self):
        return self.__class__.__name__

    def forward(self, input: SparseTensor, input_glob: SparseTensor):
        assert isinstance(input, SparseTensor)
        assert isinstance(input_glob, SparseTensor)
        zeros = torch.zeros((len(input), input_glob.F.size(1)), dtype=input_glob.dtype, device=input_glob.device)
        out = MinkowskiBroadcastFunction.apply(zeros, input_glob.F, BroadcastMode.ELEMENTWISE_ADDITON,
                                               input.coordinate_map_key, input_glob.coordinate_map_key,
                                               input.coordinate_manager)
        return SparseTensor(out, coordinate_map_key=input.coordinate_map_key,
                            coordinate_manager=input.coordinate_manager)


class MinkowskiBroadcastConcatenation(MinkowskiBroadcast):
    """y_{i, b} = [x_{i, b}, g_b]"""

    def forward(self, input: SparseTensor, input_glob: SparseTensor):
        assert isinstance(input, SparseTensor)
        assert isinstance(input_glob, SparseTensor)
        broadcast = MinkowskiBroadcast.forward(self, input, input_glob)
        return SparseTensor(torch.cat((input.F, broadcast.F), dim=1), coordinate_map_key=input.coordinate_map_key,
                            coordinate_manager=input.coordinate_manager)
