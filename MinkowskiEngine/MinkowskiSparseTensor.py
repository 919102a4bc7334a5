"""`ME.SparseTensor` (reference: MinkowskiSparseTensor.py:48-783).

Construction paths kept identical to the reference:
  * `(features, coordinates[, tensor_stride, coordinate_manager])` — a new
    coordinate map is inserted (duplicates removed; default quantisation keeps
    the first occurrence of each coordinate, RANDOM_SUBSAMPLE :339-340);
  * `(features, coordinate_map_key=, coordinate_manager=)` — wraps features on
    an existing map without touching coordinates (:198-215, :278-279).
"""
import os
import warnings

import torch

from MinkowskiEngineBackend._C import (CoordinateMapKey, CoordinateMapType, GPUMemoryAllocatorType,
                                       MinkowskiAlgorithm)

from .MinkowskiCommon import StrideType, convert_to_int_list
from .MinkowskiCoordinateManager import CoordinateManager
from .MinkowskiTensor import (SparseTensorOperationMode, SparseTensorQuantizationMode, Tensor,
                              global_coordinate_manager, set_global_coordinate_manager,
                              sparse_tensor_operation_mode)


class SparseTensor(Tensor):
    __slots__ = ("_C", "_F", "_D", "coordinate_map_key", "_manager", "unique_index", "inverse_mapping",
                 "quantization_mode", "_batch_rows")

    def __init__(self, features: torch.Tensor, coordinates: torch.Tensor = None,
                 tensor_stride: StrideType = 1, coordinate_map_key: CoordinateMapKey = None,
                 coordinate_manager: CoordinateManager = None,
                 quantization_mode: SparseTensorQuantizationMode = SparseTensorQuantizationMode.RANDOM_SUBSAMPLE,
                 allocator_type: GPUMemoryAllocatorType = None, minkowski_algorithm: MinkowskiAlgorithm = None,
                 requires_grad=None, device=None):
        assert isinstance(features, torch.Tensor), "Features must be a torch.Tensor"
        assert features.ndim == 2, f"The feature should be a matrix, The input feature is an order-{features.ndim} tensor."
        assert isinstance(quantization_mode, SparseTensorQuantizationMode)
        self.quantization_mode = quantization_mode
        if coordinates is not None:
            assert isinstance(coordinates, torch.Tensor)
        if coordinate_map_key is not None:
            assert isinstance(coordinate_map_key, CoordinateMapKey)
            assert coordinate_manager is not None, "Must provide coordinate_manager if coordinate_map_key is provided"
            assert coordinates is None, "Must not provide coordinates if coordinate_map_key is provided"
        if coordinate_manager is not None:
            assert isinstance(coordinate_manager, CoordinateManager)
        if coordinates is None and (coordinate_map_key is None or coordinate_manager is None):
            raise ValueError("Either coordinates or (coordinate_map_key, coordinate_manager) pair must be provided.")

        if device is not None:
            features = features.to(device)
            if coordinates is not None:
                coordinates = coordinates.to(device)
        self._D = coordinates.size(1) - 1 if coordinates is not None else coordinate_manager.D
        self._C = None
        self._batch_rows = None
        self.unique_index = None
        self.inverse_mapping = None

        if coordinate_manager is None:
            def _new_manager():
                return CoordinateManager(
                    D=self._D,
                    coordinate_map_type=CoordinateMapType.CUDA if coordinates.is_cuda else CoordinateMapType.CPU,
                    allocator_type=allocator_type, minkowski_algorithm=minkowski_algorithm)
            if sparse_tensor_operation_mode() == SparseTensorOperationMode.SHARE_COORDINATE_MANAGER:
                coordinate_manager = global_coordinate_manager()
                if coordinate_manager is None:
                    coordinate_manager = _new_manager()
                    set_global_coordinate_manager(coordinate_manager)
            else:
                coordinate_manager = _new_manager()
        self._manager = coordinate_manager

        if coordinates is not None:
            assert features.shape[0] == coordinates.shape[0], \
                "The number of rows in features and coordinates must match."
            assert features.is_cuda == coordinates.is_cuda, "Features and coordinates must have the same backend."
            coordinate_map_key = CoordinateMapKey(convert_to_int_list(tensor_stride, self._D), "")
            coordinates, features, coordinate_map_key = self.initialize_coordinates(
                coordinates, features, coordinate_map_key)
        else:
            assert coordinate_map_key.is_key_set(), "The coordinate key must be valid."

        if requires_grad is not None:
            features.requires_grad_(requires_grad)
        self._F = features
        self._C = coordinates
        self.coordinate_map_key = coordinate_map_key

    @property
    def coordinate_key(self):
        return self.coordinate_map_key

    def initialize_coordinates(self, coordinates, features, coordinate_map_key):
        if coordinates.dtype != torch.int32:
            warnings.warn("coordinates implicitly converted to torch.IntTensor. To remove this warning, use "
                          "`.int()` to convert the coords into an torch.IntTensor")
            coordinates = torch.floor(coordinates).int()
        coordinate_map_key, (unique_index, inverse_mapping) = self._manager.insert_and_map(
            coordinates, *coordinate_map_key.get_key())
        self.unique_index = unique_index.long()
        self.inverse_mapping = inverse_mapping
        n_in, n_u = coordinates.shape[0], self.unique_index.shape[0]
        mode = self.quantization_mode
        if n_u == n_in and mode in (SparseTensorQuantizationMode.RANDOM_SUBSAMPLE,
                                    SparseTensorQuantizationMode.NO_QUANTIZATION):
            # all rows unique: unique_index is the identity (first-occurrence order, ascending)
            return coordinates, features, coordinate_map_key
        coordinates = coordinates[self.unique_index]
        if mode in (SparseTensorQuantizationMode.UNWEIGHTED_SUM, SparseTensorQuantizationMode.UNWEIGHTED_AVERAGE):
            summed = torch.zeros((n_u, features.shape[1]), dtype=features.dtype, device=features.device)
            summed.index_add_(0, self.inverse_mapping, features)
            if mode == SparseTensorQuantizationMode.UNWEIGHTED_AVERAGE:
                cnt = torch.zeros(n_u, dtype=features.dtype, device=features.device)
                cnt.index_add_(0, self.inverse_mapping, torch.ones(n_in, dtype=features.dtype, device=features.device))
                summed = summed / cnt.unsqueeze(1)
            features = summed
        elif mode == SparseTensorQuantizationMode.RANDOM_SUBSAMPLE:
            features = features[self.unique_index]
        return coordinates, features, coordinate_map_key

    # -- dense / sparse conversion -------------------------------------------------
    def dense(self, shape=None, min_coordinate=None, contract_stride=True):
        """-> (dense [B, C, *spatial], min_coordinate, tensor_stride); the
        element at dense index x holds coordinate min_coordinate + tensor_stride * x."""
        if min_coordinate is not None:
            assert isinstance(min_coordinate, torch.IntTensor)
            assert min_coordinate.numel() == self._D
        if shape is not None:
            assert isinstance(shape, torch.Size)
            assert len(shape) == self._D + 2
            if shape[1] != self._F.size(1):
                shape = torch.Size([shape[0], self._F.size(1), *[s for s in shape[2:]]])
        if len(self) == 0:
            assert shape is not None, "shape is required to densify an empty tensor"
            return (torch.zeros(shape, dtype=self.dtype, device=self.device),
                    torch.zeros(self._D, dtype=torch.int32, device=self.device), self.tensor_stride)
        tensor_stride = torch.IntTensor(self.tensor_stride).to(self.device)
        batch_indices = self.C[:, 0]
        if min_coordinate is None:
            min_coordinate, _ = self.C.min(0, keepdim=True)
            min_coordinate = min_coordinate[:, 1:]
            if not torch.all(min_coordinate >= 0):
                raise ValueError(f"Coordinate has a negative value: {min_coordinate}. Please provide min_coordinate argument")
            coords = self.C[:, 1:]
        elif isinstance(min_coordinate, int) and min_coordinate == 0:
            coords = self.C[:, 1:]
        else:
            min_coordinate = min_coordinate.to(self.device)
            if min_coordinate.ndim == 1:
                min_coordinate = min_coordinate.unsqueeze(0)
            coords = self.C[:, 1:] - min_coordinate
        assert (min_coordinate % tensor_stride).sum() == 0, \
            "The minimum coordinates must be divisible by the tensor stride."
        if coords.ndim == 1:
            coords = coords.unsqueeze(1)
        if contract_stride:
            coords = coords // tensor_stride
        nchannels = self.F.size(1)
        if shape is None:
            size = coords.max(0)[0] + 1
            shape = torch.Size([int(batch_indices.max()) + 1, nchannels, *size.cpu().numpy()])
        dense_F = torch.zeros(shape, dtype=self.dtype, device=self.device)
        idx = (batch_indices.long(), slice(None)) + tuple(coords[:, d].long() for d in range(self._D))
        # advanced indexing puts the indexed dims first: result rows are [N, C]
        dense_F[idx[0], :, *idx[2:]] = self.F
        tensor_stride = torch.IntTensor(self.tensor_stride)
        return dense_F, min_coordinate, tensor_stride

    def sparse(self, min_coords=None, max_coords=None, contract_coords=True):
        """torch sparse COO view `[B, *spatial, C]` -> (sparse, min_coords, tensor_stride)."""
        tensor_stride = torch.IntTensor(self.tensor_stride).to(self.device)
        coords, batch_indices = self.C[:, 1:], self.C[:, 0:1]
        if min_coords is None:
            min_coords, _ = coords.min(0, keepdim=True)
        else:
            min_coords = min_coords.to(self.device).view(1, -1)
        coords = coords - min_coords
        if contract_coords:
            coords = coords // tensor_stride
            min_coords = min_coords // tensor_stride
        idx = torch.cat((batch_indices, coords), dim=1).long().t()
        size = None
        if max_coords is not None:
            mc = max_coords.to(self.device).view(1, -1)
            if contract_coords:
                mc = mc // tensor_stride
            extent = (mc - min_coords + 1).view(-1).tolist()
            size = torch.Size([int(batch_indices.max()) + 1, *[int(e) for e in extent], self.F.size(1)])
        st = torch.sparse_coo_tensor(idx, self.F, size) if size is not None else torch.sparse_coo_tensor(idx, self.F)
        return st, min_coords, torch.IntTensor(self.tensor_stride)

    def features_at_coordinates(self, query_coordinates: torch.Tensor):
        raise NotImplementedError("interpolation is outside this engine's scope (SURVEY §2.1 row 5)")

    def __repr__(self):
        return (self.__class__.__name__ + "(" + os.linesep + "  coordinates=" + str(self.C) + os.linesep
                + "  features=" + str(self.F) + os.linesep + "  coordinate_map_key=" + str(self.coordinate_map_key)
                + os.linesep + "  coordinate_manager=" + str(self._manager) + "  spatial dimension="
                + str(self._D) + ")")


def _get_coordinate_map_key(input: SparseTensor, coordinates=None, tensor_stride: StrideType = 1,
                            expand_coordinates: bool = False):
    """Output key of an op (reference :754-783): an unset key when no target is
    given (the engine fills it in), otherwise the key of the target tensor /
    map / freshly inserted coordinates.  NB callers pass `expand_coordinates`
    positionally into the `tensor_stride` slot (MinkowskiConvolution.py:319-321);
    that quirk is tolerated here: a bool tensor_stride means stride 1."""
    if isinstance(tensor_stride, bool):
        tensor_stride = 1
    if coordinates is not None and not expand_coordinates:
        assert isinstance(coordinates, (CoordinateMapKey, torch.Tensor, SparseTensor))
        if isinstance(coordinates, torch.Tensor):
            assert coordinates.ndim == 2
            key = CoordinateMapKey(convert_to_int_list(tensor_stride, coordinates.size(1) - 1), "")
            key, _ = input._manager.insert_and_map(coordinates, *key.get_key())
            return key
        if isinstance(coordinates, SparseTensor):
            return coordinates.coordinate_map_key
        return coordinates
    return CoordinateMapKey(input.coordinate_map_key.get_coordinate_size())
