"""`ME.CoordinateManager` — thin Python owner of the backend manager.

Public surface follows the reference's MinkowskiCoordinateManager.py
(:46-498): module-level allocator / map-type / algorithm switches and a
`CoordinateManager` class that forwards to `_C.CoordinateMapManager*`.
"""
import os
import warnings
from collections.abc import Sequence
from typing import List, Tuple, Union

import numpy as np
import torch

import MinkowskiEngineBackend._C as _C
from MinkowskiEngineBackend._C import (CoordinateMapKey, CoordinateMapType, GPUMemoryAllocatorType,
                                       MinkowskiAlgorithm, RegionType)

from .MinkowskiCommon import convert_to_int_list

CPU_COUNT = int(os.environ.get("OMP_NUM_THREADS", os.cpu_count() or 1))

_allocator_type = GPUMemoryAllocatorType.PYTORCH
_coordinate_map_type = CoordinateMapType.CUDA if _C.is_cuda_available() else CoordinateMapType.CPU
_minkowski_algorithm = MinkowskiAlgorithm.DEFAULT


def set_coordinate_map_type(coordinate_map_type: CoordinateMapType):
    global _coordinate_map_type
    _coordinate_map_type = coordinate_map_type


def set_gpu_allocator(backend: GPUMemoryAllocatorType):
    """Kept for API compatibility: this engine always allocates through the
    PyTorch caching allocator, both enum values are accepted."""
    assert isinstance(backend, GPUMemoryAllocatorType), \
        f"Input must be an instance of GPUMemoryAllocatorType not {backend}"
    global _allocator_type
    _allocator_type = backend


def set_memory_manager_backend(backend: GPUMemoryAllocatorType):
    warnings.warn("`set_memory_manager_backend` has been deprecated. Use `set_gpu_allocator` instead.")
    set_gpu_allocator(backend)


class CoordsManager:
    def __init__(*args, **kwargs):
        raise DeprecationWarning("`CoordsManager` has been deprecated. Use `CoordinateManager` instead.")


class CoordinateManager:
    def __init__(self, D: int = 0, num_threads: int = -1, coordinate_map_type: CoordinateMapType = None,
                 allocator_type: GPUMemoryAllocatorType = None, minkowski_algorithm: MinkowskiAlgorithm = None):
        if D < 1:
            raise ValueError(f"Invalid rank D > 0, D = {D}.")
        if num_threads < 0:
            num_threads = min(CPU_COUNT, 20)
        coordinate_map_type = _coordinate_map_type if coordinate_map_type is None else coordinate_map_type
        allocator_type = _allocator_type if allocator_type is None else allocator_type
        minkowski_algorithm = _minkowski_algorithm if minkowski_algorithm is None else minkowski_algorithm
        if coordinate_map_type == CoordinateMapType.CPU:
            postfix = "CPU"
        else:
            postfix = "GPU" + ("_default" if allocator_type == GPUMemoryAllocatorType.CUDA else "_c10")
        self.D = D
        self.minkowski_algorithm = minkowski_algorithm
        self._CoordinateManagerClass = getattr(_C, "CoordinateMapManager" + postfix)
        self._manager = self._CoordinateManagerClass(minkowski_algorithm, num_threads)

    # -- map creation ---------------------------------------------------------
    def insert_and_map(self, coordinates: torch.Tensor, tensor_stride: Union[int, Sequence, np.ndarray] = 1,
                       string_id: str = "") -> Tuple[CoordinateMapKey, Tuple[torch.Tensor, torch.Tensor]]:
        """-> (key, (unique_map, inverse_map)) with
        coordinates[unique_map] == get_coordinates(key) and
        coordinates == coordinates[unique_map][inverse_map]."""
        return self._manager.insert_and_map(coordinates, convert_to_int_list(tensor_stride, self.D), string_id)

    def stride(self, coordinate_map_key: CoordinateMapKey, stride, string_id: str = "") -> CoordinateMapKey:
        return self._manager.stride(coordinate_map_key, convert_to_int_list(stride, self.D), string_id)

    def origin(self) -> CoordinateMapKey:
        return self._manager.origin()

    def size(self, coordinate_map_key: CoordinateMapKey) -> int:
        return self._manager.size(coordinate_map_key)

    # -- lookup -----------------------------------------------------------------
    def _get_coordinate_map_key(self, key_or_tensor_strides) -> CoordinateMapKey:
        assert isinstance(key_or_tensor_strides, (CoordinateMapKey, Sequence, np.ndarray, torch.Tensor, int)), \
            f"The input must be either a CoordinateMapKey or tensor_stride. Invalid: {key_or_tensor_strides}"
        if isinstance(key_or_tensor_strides, CoordinateMapKey):
            return key_or_tensor_strides
        keys = self._manager.get_coordinate_map_keys(convert_to_int_list(key_or_tensor_strides, self.D))
        assert len(keys) > 0
        return keys[0]

    def get_coordinates(self, coords_key_or_tensor_strides) -> torch.Tensor:
        return self._manager.get_coordinates(self._get_coordinate_map_key(coords_key_or_tensor_strides))

    def number_of_unique_batch_indices(self) -> int:
        return self._manager.origin_map_size()

    def get_unique_coordinate_map_key(self, tensor_stride: Union[int, list]) -> CoordinateMapKey:
        return self._manager.get_random_string_id(convert_to_int_list(tensor_stride, self.D), "")

    # -- maps ---------------------------------------------------------------------
    def kernel_map(self, in_key, out_key, stride=1, kernel_size=3, dilation=1,
                   region_type=RegionType.HYPER_CUBE, region_offset=None, is_transpose=False,
                   is_pool=False) -> dict:
        """dict{kernel_index: in_out_tensor}; in_out_tensor[0] = input rows,
        in_out_tensor[1] = output rows."""
        if isinstance(kernel_size, torch.Tensor):
            assert (kernel_size > 0).all(), f"Invalid kernel size: {kernel_size}"
            if (kernel_size == 1).all():
                region_type = RegionType.HYPER_CUBE
        elif isinstance(kernel_size, int):
            assert kernel_size > 0, f"Invalid kernel size: {kernel_size}"
            if kernel_size == 1:
                region_type = RegionType.HYPER_CUBE
        in_key = self._get_coordinate_map_key(in_key)
        out_key = self._get_coordinate_map_key(out_key)
        if region_offset is None:
            region_offset = torch.IntTensor()
        return self._manager.kernel_map(in_key, out_key, convert_to_int_list(kernel_size, self.D),
                                        convert_to_int_list(stride, self.D), convert_to_int_list(dilation, self.D),
                                        region_type, region_offset, is_transpose, is_pool)

    def get_kernel_map(self, *args, **kwargs) -> dict:
        warnings.warn("`get_kernel_map` will be deprecated. Please use `kernel_map` instead.")
        return self.kernel_map(*args, **kwargs)

    def origin_map(self, key: CoordinateMapKey):
        return self._manager.origin_map(key)

    def stride_map(self, in_key: CoordinateMapKey, stride_key: CoordinateMapKey):
        return self._manager.stride_map(in_key, stride_key)

    def union_map(self, in_keys: list, out_key):
        return self._manager.union_map(in_keys, out_key)

    # -- unsupported continuous-coordinate API (out of scope) ---
    def _no_field(self, *args, **kwargs):
        raise NotImplementedError("TensorField / interpolation maps are outside this engine's scope")

    insert_field = field_to_sparse_insert_and_map = exists_field_to_sparse = field_to_sparse_keys = _no_field
    get_field_to_sparse_map = field_to_sparse_map = origin_field = get_coordinate_field = _no_field
    origin_field_map = interpolation_map_weight = _no_field

    def __repr__(self):
        return (self._CoordinateManagerClass.__name__ + "(\n" + str(self._manager)
                + f"\talgorithm={self.minkowski_algorithm}\n  )\n")
