"""Kernel-shape bookkeeping (host only).  Same public names as the reference's
MinkowskiKernelGenerator.py (KernelGenerator, KernelRegion, get_kernel_volume,
save_ctx)."""
from collections import namedtuple
from functools import reduce

import torch

from MinkowskiEngineBackend._C import RegionType

from .MinkowskiCommon import convert_to_int_list


def get_kernel_volume(region_type, kernel_size, region_offset, axis_types, dimension):
    """HYPER_CUBE: prod(kernel_size); HYPER_CROSS: sum(k-1)+1 (odd sizes);
    CUSTOM: number of offsets (reference :36-102)."""
    ks = [int(k) for k in kernel_size]
    if region_type == RegionType.HYPER_CUBE:
        assert all(k > 0 for k in ks), "kernel_size must be positive"
        assert region_offset is None or len(region_offset) == 0, \
            "Region offset must be None when region_type is given"
        assert axis_types is None, "Axis types must be None when region_type is given"
        return int(reduce(lambda a, b: a * b, ks, 1))
    if region_type == RegionType.HYPER_CROSS:
        assert all(k > 0 for k in ks), "kernel_size must be positive"
        assert all(k % 2 == 1 for k in ks), "kernel_size must be odd for region_type HYPER_CROSS"
        return int(sum(k - 1 for k in ks) + 1)
    if region_type == RegionType.CUSTOM:
        assert region_offset is not None and region_offset.numel() > 0, \
            "region_offset must be non empty when region_type is CUSTOM"
        assert region_offset.size(1) == dimension, \
            "region_offset must have the same dimension as the network"
        return int(region_offset.size(0))
    raise NotImplementedError()


class KernelGenerator:
    __slots__ = ("cache", "kernel_size", "kernel_stride", "kernel_dilation", "region_type", "region_offsets",
                 "axis_types", "dimension", "kernel_volume", "requires_strided_coordinates", "expand_coordinates")

    def __init__(self, kernel_size=-1, stride=1, dilation=1, is_transpose: bool = False,
                 region_type: RegionType = RegionType.HYPER_CUBE, region_offsets: torch.Tensor = None,
                 expand_coordinates: bool = False, axis_types=None, dimension=-1):
        assert dimension > 0
        assert isinstance(region_type, RegionType)
        self.cache = {}
        self.kernel_size = convert_to_int_list(kernel_size, dimension)
        self.kernel_stride = convert_to_int_list(stride, dimension)
        self.kernel_dilation = convert_to_int_list(dilation, dimension)
        self.region_type = region_type
        self.region_offsets = region_offsets if region_offsets is not None and region_offsets.numel() > 0 \
            else torch.IntTensor()
        self.axis_types = axis_types
        self.dimension = dimension
        self.kernel_volume = get_kernel_volume(region_type, self.kernel_size, region_offsets, axis_types, dimension)
        # NB (reference :307-309): True when every stride == 1; gates the use_mm shortcut
        self.requires_strided_coordinates = all(s == 1 for s in self.kernel_stride)
        self.expand_coordinates = expand_coordinates

    def __repr__(self):
        return (self.__class__.__name__ + f"(kernel_size={self.kernel_size}, kernel_stride={self.kernel_stride}, "
                f"kernel_dilation={self.kernel_dilation}, region_type={self.region_type}, "
                f"expand_coordinates={self.expand_coordinates}, dimension={self.dimension})")


class KernelRegion(namedtuple("KernelRegion", ("kernel_size", "kernel_stride", "kernel_dilation", "region_type",
                                               "offset", "D"))):
    """adding functionality to a named tuple"""
    __slots__ = ()

    def __init__(self, kernel_size, kernel_stride, kernel_dilation, region_type, offset, dimension):
        kernel_size = convert_to_int_list(kernel_size, dimension)
        kernel_stride = convert_to_int_list(kernel_stride, dimension)
        kernel_dilation = convert_to_int_list(kernel_dilation, dimension)
        super(KernelRegion, self).__init__(kernel_size, kernel_stride, kernel_dilation, region_type, offset,
                                           dimension)

    def __str__(self):
        return "kernel_size:{self.kernel_size}, kernel_stride:{self.kernel_stride}, region_type:{self.region_type}"


def save_ctx(ctx, kernel_generator: KernelGenerator, in_coords_key, out_coords_key, coordinate_manager):
    ctx.kernel_generator = kernel_generator
    ctx.in_coordinate_map_key = in_coords_key
    ctx.out_coordinate_map_key = out_coords_key
    ctx.coordinate_manager = coordinate_manager
    return ctx
