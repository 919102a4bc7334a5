"""Sparse convolution layers of the ME API on the B200 engine.

Public classes and their constructor/forward signatures follow the reference
(MinkowskiConvolution.py:42-736): the two autograd Functions, the shared
module base (kernel `[K, Cin, Cout]`, or `[Cin, Cout]` when the kernel volume
is 1 at stride 1 — the `use_mm` shortcut :312-316), `MinkowskiConvolution`,
`MinkowskiConvolutionTranspose`, `MinkowskiGenerativeConvolutionTranspose` and
the fork's `MinkowskiUpsampleInterpolate` (:644-735, frozen identity kernels).
Parameter names (`kernel`, `bias`) and shapes are unchanged, so reference
checkpoints load.
"""
import math
from typing import Union

import torch
from torch.autograd import Function
from torch.nn import Parameter

from MinkowskiEngineBackend._C import ConvolutionMode, CoordinateMapKey, RegionType

from .MinkowskiCommon import MinkowskiModuleBase, get_minkowski_function
from .MinkowskiCoordinateManager import CoordinateManager
from .MinkowskiKernelGenerator import KernelGenerator
from .MinkowskiSparseTensor import SparseTensor, _get_coordinate_map_key


def _match_dtype(features: torch.Tensor, kernel: torch.Tensor) -> torch.Tensor:
    """The engine wants equal scalar types (reference assert).  Under autocast /
    a bf16 model the features decide: the kernel is cast to the feature dtype."""
    return kernel if kernel.dtype == features.dtype else kernel.to(features.dtype)


def _autocast(features: torch.Tensor) -> torch.Tensor:
    """Under `torch.autocast("cuda", dtype=bf16|f16)` (Lightning's `precision="bf16-mixed"`,
    examples/ae_res.py:1016) the convolution is a matmul-class op: features (and, through
    `_match_dtype`, the kernel) are cast to the autocast dtype, accumulation stays fp32."""
    if features.is_cuda and features.is_floating_point() and torch.is_autocast_enabled("cuda"):
        dt = torch.get_autocast_dtype("cuda")
        if features.dtype != dt and features.dtype != torch.float64:
            return features.to(dt)
    return features


class _ConvFunctionBase(Function):
    FWD = BWD = None

    @classmethod
    def _forward(cls, ctx, input_features, kernel_weights, kernel_generator, convolution_mode,
                 in_coordinate_map_key, out_coordinate_map_key, coordinate_manager):
        if out_coordinate_map_key is None:
            out_coordinate_map_key = CoordinateMapKey(in_coordinate_map_key.get_coordinate_size())
        input_features = input_features.contiguous()
        ctx.input_features = input_features
        ctx.kernel_weights = kernel_weights
        ctx.misc = [kernel_generator, convolution_mode, in_coordinate_map_key, out_coordinate_map_key,
                    coordinate_manager]
        fw_fn = get_minkowski_function(cls.FWD, input_features)
        return fw_fn(input_features, _match_dtype(input_features, kernel_weights), kernel_generator.kernel_size,
                     kernel_generator.kernel_stride, kernel_generator.kernel_dilation,
                     kernel_generator.region_type, kernel_generator.region_offsets,
                     kernel_generator.expand_coordinates, convolution_mode, in_coordinate_map_key,
                     out_coordinate_map_key, coordinate_manager._manager)

    @classmethod
    def _backward(cls, ctx, grad_out_feat):
        grad_out_feat = grad_out_feat.contiguous()
        kernel_generator, convolution_mode, in_key, out_key, coordinate_manager = ctx.misc
        x = ctx.input_features
        if grad_out_feat.dtype != x.dtype:
            grad_out_feat = grad_out_feat.to(x.dtype)
        bw_fn = get_minkowski_function(cls.BWD, grad_out_feat)
        grad_in_feat, grad_kernel = bw_fn(x, grad_out_feat, _match_dtype(x, ctx.kernel_weights),
                                          kernel_generator.kernel_size, kernel_generator.kernel_stride,
                                          kernel_generator.kernel_dilation, kernel_generator.region_type,
                                          kernel_generator.region_offsets, convolution_mode, in_key, out_key,
                                          coordinate_manager._manager)
        if grad_kernel.dtype != ctx.kernel_weights.dtype:
            grad_kernel = grad_kernel.to(ctx.kernel_weights.dtype)
        return grad_in_feat, grad_kernel, None, None, None, None, None


class MinkowskiConvolutionFunction(_ConvFunctionBase):
    FWD, BWD = "ConvolutionForward", "ConvolutionBackward"

    @staticmethod
    def forward(ctx, input_features: torch.Tensor, kernel_weights: torch.Tensor,
                kernel_generator: KernelGenerator, convolution_mode: ConvolutionMode,
                in_coordinate_map_key: CoordinateMapKey, out_coordinate_map_key: CoordinateMapKey = None,
                coordinate_manager: CoordinateManager = None):
        return MinkowskiConvolutionFunction._forward(ctx, input_features, kernel_weights, kernel_generator,
                                                     convolution_mode, in_coordinate_map_key,
                                                     out_coordinate_map_key, coordinate_manager)

    @staticmethod
    def backward(ctx, grad_out_feat: torch.Tensor):
        return MinkowskiConvolutionFunction._backward(ctx, grad_out_feat)


class MinkowskiConvolutionTransposeFunction(_ConvFunctionBase):
    FWD, BWD = "ConvolutionTransposeForward", "ConvolutionTransposeBackward"

    @staticmethod
    def forward(ctx, input_features: torch.Tensor, kernel_weights: torch.Tensor,
                kernel_generator: KernelGenerator, convolution_mode: ConvolutionMode,
                in_coordinate_map_key: CoordinateMapKey, out_coordinate_map_key: CoordinateMapKey = None,
                coordinate_manager: CoordinateManager = None):
        return MinkowskiConvolutionTransposeFunction._forward(ctx, input_features, kernel_weights,
                                                              kernel_generator, convolution_mode,
                                                              in_coordinate_map_key, out_coordinate_map_key,
                                                              coordinate_manager)

    @staticmethod
    def backward(ctx, grad_out_feat: torch.Tensor):
        return MinkowskiConvolutionTransposeFunction._backward(ctx, grad_out_feat)


class MinkowskiConvolutionBase(MinkowskiModuleBase):

    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, is_transpose=False, is_interpolation=False, expand_coordinates=False,
                 convolution_mode=ConvolutionMode.DEFAULT, dimension=-1):
        super().__init__()
        assert dimension > 0, f"Invalid dimension. Please provide a valid dimension argument. dimension={dimension}"
        if kernel_generator is None:
            kernel_generator = KernelGenerator(kernel_size=kernel_size, stride=stride, dilation=dilation,
                                               expand_coordinates=expand_coordinates, dimension=dimension)
        else:
            kernel_generator.expand_coordinates = expand_coordinates
        self.is_transpose = is_transpose
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.kernel_generator = kernel_generator
        self.dimension = dimension
        K = kernel_generator.kernel_volume
        # kernel volume 1 at stride 1 == plain matrix product on the features
        self.use_mm = K == 1 and kernel_generator.requires_strided_coordinates
        shape = (in_channels, out_channels) if self.use_mm else (K, in_channels, out_channels)
        if is_interpolation:
            assert in_channels == out_channels
            self.kernel = Parameter(torch.eye(out_channels).repeat(K, 1, 1), requires_grad=False)
        else:
            self.kernel = Parameter(torch.empty(*shape, dtype=torch.float32))
        self.bias = Parameter(torch.empty(1, out_channels, dtype=torch.float32)) if bias else None
        self.convolution_mode = convolution_mode
        self.conv = MinkowskiConvolutionTransposeFunction if is_transpose else MinkowskiConvolutionFunction

    def forward(self, input: SparseTensor,
                coordinates: Union[torch.Tensor, CoordinateMapKey, SparseTensor] = None):
        assert isinstance(input, SparseTensor)
        assert input.D == self.dimension
        if self.use_mm:
            out_key = input.coordinate_map_key
            outfeat = input.F.mm(_match_dtype(input.F, self.kernel))
        else:
            out_key = _get_coordinate_map_key(input, coordinates, self.kernel_generator.expand_coordinates)
            outfeat = self.conv.apply(_autocast(input.F), self.kernel, self.kernel_generator, self.convolution_mode,
                                      input.coordinate_map_key, out_key, input._manager)
        if self.bias is not None:
            outfeat = outfeat + _match_dtype(outfeat, self.bias)
        return SparseTensor(outfeat, coordinate_map_key=out_key, coordinate_manager=input._manager)

    def reset_parameters(self, is_transpose=False):
        with torch.no_grad():
            n = (self.out_channels if is_transpose else self.in_channels) * self.kernel_generator.kernel_volume
            stdv = 1.0 / math.sqrt(n)
            self.kernel.data.uniform_(-stdv, stdv)
            if self.bias is not None:
                self.bias.data.uniform_(-stdv, stdv)

    def __repr__(self):
        kg = self.kernel_generator
        s = f"(in={self.in_channels}, out={self.out_channels}, "
        if kg.region_type in [RegionType.CUSTOM]:
            s += f"region_type={kg.region_type}, kernel_volume={kg.kernel_volume}, "
        else:
            s += f"kernel_size={kg.kernel_size}, "
        s += f"stride={kg.kernel_stride}, dilation={kg.kernel_dilation})"
        return self.__class__.__name__ + s


class MinkowskiConvolution(MinkowskiConvolutionBase):
    """x_u = sum_{i in N(u, K, C_in)} W_i x_{i+u}   (generalised sparse convolution)."""

    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, expand_coordinates=False, convolution_mode=ConvolutionMode.DEFAULT,
                 dimension=None):
        MinkowskiConvolutionBase.__init__(self, in_channels, out_channels, kernel_size, stride, dilation, bias,
                                          kernel_generator, is_transpose=False,
                                          expand_coordinates=expand_coordinates,
                                          convolution_mode=convolution_mode, dimension=dimension)
        self.reset_parameters()


class MinkowskiConvolutionTranspose(MinkowskiConvolutionBase):
    """Transposed sparse convolution: output tensor stride = input / stride."""

    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, expand_coordinates=False, convolution_mode=ConvolutionMode.DEFAULT,
                 dimension=None):
        if kernel_generator is None:
            kernel_generator = KernelGenerator(kernel_size=kernel_size, stride=stride, dilation=dilation,
                                               dimension=dimension)
        MinkowskiConvolutionBase.__init__(self, in_channels, out_channels, kernel_size, stride, dilation, bias,
                                          kernel_generator, is_transpose=True,
                                          expand_coordinates=expand_coordinates,
                                          convolution_mode=convolution_mode, dimension=dimension)
        self.reset_parameters(True)


class MinkowskiGenerativeConvolutionTranspose(MinkowskiConvolutionBase):
    """Transposed convolution that creates new coordinates (octree split for
    kernel 2 / stride 2): expand_coordinates is forced on."""

    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, convolution_mode=ConvolutionMode.DEFAULT, dimension=None):
        if kernel_generator is None:
            kernel_generator = KernelGenerator(kernel_size=kernel_size, stride=stride, dilation=dilation,
                                               expand_coordinates=True, dimension=dimension)
        else:
            kernel_generator.expand_coordinates = True
        MinkowskiConvolutionBase.__init__(self, in_channels, out_channels, kernel_size, stride, dilation, bias,
                                          kernel_generator, is_transpose=True, expand_coordinates=True,
                                          convolution_mode=convolution_mode, dimension=dimension)
        self.reset_parameters(True)


class MinkowskiUpsampleInterpolate(MinkowskiConvolutionBase):
    """Fork addition: generative transposed convolution with frozen identity
    kernels (nearest-neighbour upsampling onto new coordinates)."""

    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, convolution_mode=ConvolutionMode.DEFAULT, dimension=None):
        if kernel_generator is None:
            kernel_generator = KernelGenerator(kernel_size=kernel_size, stride=stride, dilation=dilation,
                                               expand_coordinates=True, dimension=dimension)
        else:
            kernel_generator.expand_coordinates = True
        MinkowskiConvolutionBase.__init__(self, in_channels, out_channels, kernel_size, stride, dilation, bias,
                                          kernel_generator, is_transpose=True, is_interpolation=True,
                                          expand_coordinates=True, convolution_mode=convolution_mode,
                                          dimension=dimension)
