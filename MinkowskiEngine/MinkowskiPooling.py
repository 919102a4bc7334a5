"""Pooling layers (reference: MinkowskiPooling.py:42-780).

Local pooling reduces over the same kernel map a convolution would use
(sum / average over the existing neighbours / max); global pooling reduces
every batch item to one row of the origin map.
"""
from typing import Union

import torch
from torch.autograd import Function

from MinkowskiEngineBackend._C import CoordinateMapKey, PoolingMode

from .MinkowskiCommon import MinkowskiModuleBase, get_minkowski_function
from .MinkowskiCoordinateManager import CoordinateManager
from .MinkowskiKernelGenerator import KernelGenerator
from .MinkowskiSparseTensor import SparseTensor, _get_coordinate_map_key


class _LocalPoolFn(Function):
    FWD = BWD = None
    TRANSPOSE = False

    @classmethod
    def _forward(cls, ctx, input_features, pooling_mode, kernel_generator, in_key, out_key, coordinate_manager):
        if out_key is None:
            out_key = CoordinateMapKey(in_key.get_coordinate_size())
        input_features = input_features.contiguous()
        ctx.input_features = input_features
        ctx.misc = (pooling_mode, kernel_generator, in_key, out_key, coordinate_manager)
        fw_fn = get_minkowski_function(cls.FWD, input_features)
        args = [input_features, kernel_generator.kernel_size, kernel_generator.kernel_stride,
                kernel_generator.kernel_dilation, kernel_generator.region_type, kernel_generator.region_offsets]
        if cls.TRANSPOSE:
            args.append(kernel_generator.expand_coordinates)
        out_feat, num_nonzero = fw_fn(*args, pooling_mode, in_key, out_key, coordinate_manager._manager)
        ctx.num_nonzero = num_nonzero
        return out_feat

    @classmethod
    def _backward(cls, ctx, grad_out_feat):
        grad_out_feat = grad_out_feat.contiguous()
        pooling_mode, kernel_generator, in_key, out_key, coordinate_manager = ctx.misc
        bw_fn = get_minkowski_function(cls.BWD, grad_out_feat)
        grad_in_feat = bw_fn(ctx.input_features, grad_out_feat, ctx.num_nonzero, kernel_generator.kernel_size,
                             kernel_generator.kernel_stride, kernel_generator.kernel_dilation,
                             kernel_generator.region_type, kernel_generator.region_offsets, pooling_mode,
                             in_key, out_key, coordinate_manager._manager)
        return grad_in_feat, None, None, None, None, None


class MinkowskiLocalPoolingFunction(_LocalPoolFn):
    FWD, BWD = "LocalPoolingForward", "LocalPoolingBackward"

    @staticmethod
    def forward(ctx, input_features: torch.Tensor, pooling_mode: PoolingMode, kernel_generator: KernelGenerator,
                in_coordinate_map_key: CoordinateMapKey, out_coordinate_map_key: CoordinateMapKey = None,
                coordinate_manager: CoordinateManager = None):
        return MinkowskiLocalPoolingFunction._forward(ctx, input_features, pooling_mode, kernel_generator,
                                                      in_coordinate_map_key, out_coordinate_map_key,
                                                      coordinate_manager)

    @staticmethod
    def backward(ctx, grad_out_feat):
        return MinkowskiLocalPoolingFunction._backward(ctx, grad_out_feat)


class MinkowskiLocalPoolingTransposeFunction(_LocalPoolFn):
    FWD, BWD = "LocalPoolingTransposeForward", "LocalPoolingTransposeBackward"
    TRANSPOSE = True

    @staticmethod
    def forward(ctx, input_features: torch.Tensor, pooling_mode: PoolingMode, kernel_generator: KernelGenerator,
                in_coordinate_map_key: CoordinateMapKey, out_coordinate_map_key: CoordinateMapKey = None,
                coordinate_manager: CoordinateManager = None):
        return MinkowskiLocalPoolingTransposeFunction._forward(ctx, input_features, pooling_mode,
                                                               kernel_generator, in_coordinate_map_key,
                                                               out_coordinate_map_key, coordinate_manager)

    @staticmethod
    def backward(ctx, grad_out_feat):
        return MinkowskiLocalPoolingTransposeFunction._backward(ctx, grad_out_feat)


class MinkowskiPoolingBase(MinkowskiModuleBase):

    def __init__(self, kernel_size, stride=1, dilation=1, kernel_generator=None, is_transpose=False,
                 pooling_mode=PoolingMode.LOCAL_AVG_POOLING, dimension=-1):
        super().__init__()
        assert dimension > 0, f"Invalid dimension. Please provide a valid dimension argument. dimension={dimension}"
        if kernel_generator is None:
            kernel_generator = KernelGenerator(kernel_size=kernel_size, stride=stride, dilation=dilation,
                                               dimension=dimension)
        self.is_transpose = is_transpose
        self.kernel_generator = kernel_generator
        self.pooling_mode = pooling_mode
        self.dimension = dimension
        self.pooling = MinkowskiLocalPoolingFunction

    def forward(self, input: SparseTensor,
                coordinates: Union[torch.IntTensor, CoordinateMapKey, SparseTensor] = None):
        assert isinstance(input, SparseTensor)
        assert input.D == self.dimension
        out_key = _get_coordinate_map_key(input, coordinates)
        outfeat = self.pooling.apply(input.F, self.pooling_mode, self.kernel_generator, input.coordinate_map_key,
                                     out_key, input._manager)
        return SparseTensor(outfeat, coordinate_map_key=out_key, coordinate_manager=input.coordinate_manager)

    def __repr__(self):
        kg = self.kernel_generator
        return self.__class__.__name__ + \
            f"(kernel_size={kg.kernel_size}, stride={kg.kernel_stride}, dilation={kg.kernel_dilation})"


class MinkowskiAvgPooling(MinkowskiPoolingBase):
    """Average of the *existing* inputs in the kernel support (divides by the
    per-output cardinality)."""

    def __init__(self, kernel_size=-1, stride=1, dilation=1, kernel_generator=None, dimension=None):
        MinkowskiPoolingBase.__init__(self, kernel_size, stride, dilation, kernel_generator, is_transpose=False,
                                      pooling_mode=PoolingMode.LOCAL_AVG_POOLING, dimension=dimension)


class MinkowskiSumPooling(MinkowskiPoolingBase):
    def __init__(self, kernel_size=-1, stride=1, dilation=1, kernel_generator=None, dimension=None):
        MinkowskiPoolingBase.__init__(self, kernel_size, stride, dilation, kernel_generator, is_transpose=False,
                                      pooling_mode=PoolingMode.LOCAL_SUM_POOLING, dimension=dimension)


class MinkowskiMaxPooling(MinkowskiPoolingBase):
    def __init__(self, kernel_size, stride=1, dilation=1, kernel_generator=None, dimension=None):
        MinkowskiPoolingBase.__init__(self, kernel_size, stride, dilation, kernel_generator, is_transpose=False,
                                      pooling_mode=PoolingMode.LOCAL_MAX_POOLING, dimension=dimension)


class MinkowskiPoolingTranspose(MinkowskiPoolingBase):
    """Unpooling: sum of the coarse inputs that cover each fine output."""

    def __init__(self, kernel_size, stride, dilation=1, kernel_generator=None, expand_coordinates=False,
                 dimension=None):
        MinkowskiPoolingBase.__init__(self, kernel_size, stride, dilation, kernel_generator, is_transpose=True,
                                      pooling_mode=PoolingMode.LOCAL_SUM_POOLING, dimension=dimension)
        self.kernel_generator.expand_coordinates = expand_coordinates
        self.pooling = MinkowskiLocalPoolingTransposeFunction


class MinkowskiGlobalPoolingFunction(Function):
    @staticmethod
    def forward(ctx, input_features: torch.Tensor, pooling_mode: PoolingMode,
                in_coordinate_map_key: CoordinateMapKey, out_coordinate_map_key: CoordinateMapKey = None,
                coordinate_manager: CoordinateManager = None):
        if out_coordinate_map_key is None:
            out_coordinate_map_key = CoordinateMapKey(in_coordinate_map_key.get_coordinate_size())
        input_features = input_features.contiguous()
        ctx.input_features = input_features
        ctx.misc = (pooling_mode, in_coordinate_map_key, out_coordinate_map_key, coordinate_manager)
        fw_fn = get_minkowski_function("GlobalPoolingForward", input_features)
        out_feat, num_nonzero = fw_fn(input_features, pooling_mode, in_coordinate_map_key, out_coordinate_map_key,
                                      coordinate_manager._manager)
        ctx.num_nonzero = num_nonzero
        return out_feat

    @staticmethod
    def backward(ctx, grad_out_feat):
        grad_out_feat = grad_out_feat.contiguous()
        pooling_mode, in_key, out_key, coordinate_manager = ctx.misc
        bw_fn = get_minkowski_function("GlobalPoolingBackward", grad_out_feat)
        grad_in_feat = bw_fn(ctx.input_features, grad_out_feat, ctx.num_nonzero, pooling_mode, in_key, out_key,
                             coordinate_manager._manager)
        return grad_in_feat, None, None, None, None


class MinkowskiGlobalPooling(MinkowskiModuleBase):
    """Pool all input features of a batch item to one output row at the origin."""

    def __init__(self, mode: PoolingMode = PoolingMode.GLOBAL_AVG_POOLING_PYTORCH_INDEX):
        super().__init__()
        assert isinstance(mode, PoolingMode), f"Mode must be an instance of PoolingMode. mode={mode}"
        self.pooling_mode = mode
        self.pooling = MinkowskiGlobalPoolingFunction

    def forward(self, input: SparseTensor,
                coordinates: Union[torch.IntTensor, CoordinateMapKey, SparseTensor] = None):
        out_key = _get_coordinate_map_key(input, coordinates)
        output = self.pooling.apply(input.F, self.pooling_mode, input.coordinate_map_key, out_key, input._manager)
        return SparseTensor(output, coordinate_map_key=out_key, coordinate_manager=input.coordinate_manager)

    def __repr__(self):
        return self.__class__.__name__ + f"(mode={str(self.pooling_mode)})"


class MinkowskiGlobalSumPooling(MinkowskiGlobalPooling):
    def __init__(self, mode=PoolingMode.GLOBAL_SUM_POOLING_PYTORCH_INDEX):
        MinkowskiGlobalPooling.__init__(self, mode=mode)


class MinkowskiGlobalAvgPooling(MinkowskiGlobalPooling):
    def __init__(self, mode=PoolingMode.GLOBAL_AVG_POOLING_PYTORCH_INDEX):
        MinkowskiGlobalPooling.__init__(self, mode=mode)


class MinkowskiGlobalMaxPooling(MinkowskiGlobalPooling):
    def __init__(self, mode=PoolingMode.GLOBAL_MAX_POOLING_PYTORCH_INDEX):
        MinkowskiGlobalPooling.__init__(self, mode=mode)
