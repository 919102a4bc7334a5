"""Base `Tensor` of the ME API: features + a coordinate-map key + a manager.

Public names follow the reference's MinkowskiTensor.py (:39-604):
SparseTensorOperationMode, SparseTensorQuantizationMode, the global
coordinate-manager helpers and the `Tensor` base class with batch
decomposition and element-wise arithmetic (union semantics for different keys,
:511-546).
"""
from enum import Enum

import torch

from MinkowskiEngineBackend._C import CoordinateMapKey

from .MinkowskiCoordinateManager import CoordinateManager


class SparseTensorOperationMode(Enum):
    SEPARATE_COORDINATE_MANAGER = 0
    SHARE_COORDINATE_MANAGER = 1


class SparseTensorQuantizationMode(Enum):
    RANDOM_SUBSAMPLE = 0
    UNWEIGHTED_AVERAGE = 1
    UNWEIGHTED_SUM = 2
    NO_QUANTIZATION = 3
    MAX_POOL = 4
    SPLAT_LINEAR_INTERPOLATION = 5


_sparse_tensor_operation_mode = SparseTensorOperationMode.SEPARATE_COORDINATE_MANAGER
_global_coordinate_manager = None

COORDINATE_MANAGER_DIFFERENT_ERROR = "SparseTensors must share the same coordinate manager for this operation. Please refer to the SparseTensor creation API (https://nvidia.github.io/MinkowskiEngine/sparse_tensor.html) to share the coordinate manager, or set the sparse tensor operation mode with `set_sparse_tensor_operation_mode` to share it by default."
COORDINATE_KEY_DIFFERENT_ERROR = "SparseTensors must have the same coordinate_map_key."


def set_sparse_tensor_operation_mode(operation_mode: SparseTensorOperationMode):
    assert isinstance(operation_mode, SparseTensorOperationMode), \
        f"Input must be an instance of SparseTensorOperationMode not {operation_mode}"
    global _sparse_tensor_operation_mode
    _sparse_tensor_operation_mode = operation_mode


def sparse_tensor_operation_mode() -> SparseTensorOperationMode:
    return _sparse_tensor_operation_mode


def global_coordinate_manager():
    return _global_coordinate_manager


def set_global_coordinate_manager(coordinate_manager):
    global _global_coordinate_manager
    _global_coordinate_manager = coordinate_manager


def clear_global_coordinate_manager():
    set_global_coordinate_manager(None)


class Tensor:
    """Sparse tensor base: `F` [N, C] features; coordinates live in the manager
    under `coordinate_map_key` and are fetched lazily as `C` [N, D+1] int32
    (batch index first)."""

    __slots__ = ("_C", "_F", "_D", "coordinate_map_key", "_manager", "unique_index", "inverse_mapping",
                 "quantization_mode", "_batch_rows")

    @property
    def coordinate_manager(self):
        return self._manager

    @property
    def tensor_stride(self):
        return self.coordinate_map_key.get_tensor_stride()

    @tensor_stride.setter
    def tensor_stride(self, p):
        raise SyntaxError("Direct modification of tensor_stride is not permitted")

    def _get_coordinates(self):
        return self._manager.get_coordinates(self.coordinate_map_key)

    @property
    def coordinates(self):
        if self._C is None:
            self._C = self._get_coordinates()
        return self._C

    @coordinates.setter
    def coordinates(self, v):
        raise SyntaxError("Direct modification of coordinates is not permitted")

    C = coordinates

    @property
    def coordinate_key(self):
        raise NotImplementedError("Tensor interface does not have coordinate_key")

    @property
    def F(self):
        return self._F

    @property
    def features(self):
        return self._F

    # -- batch decomposition ----------------------------------------------------
    @property
    def _batchwise_row_indices(self):
        if self._batch_rows is None:
            _, self._batch_rows = self._manager.origin_map(self.coordinate_map_key)
        return self._batch_rows

    @property
    def decomposition_permutations(self):
        return self._batchwise_row_indices

    @property
    def decomposed_coordinates(self):
        return [self.C[rows, 1:] for rows in self._batchwise_row_indices]

    def coordinates_at(self, batch_index):
        return self.C[self._batchwise_row_indices[batch_index], 1:]

    @property
    def decomposed_features(self):
        return [self._F[rows] for rows in self._batchwise_row_indices]

    def features_at(self, batch_index):
        return self._F[self._batchwise_row_indices[batch_index]]

    def coordinates_and_features_at(self, batch_index):
        rows = self._batchwise_row_indices[batch_index]
        return self.C[rows, 1:], self._F[rows]

    @property
    def decomposed_coordinates_and_features(self):
        rows_list = self._batchwise_row_indices
        return [self.C[r, 1:] for r in rows_list], [self._F[r] for r in rows_list]

    # -- tensor-like attributes ---------------------------------------------------
    @property
    def dimension(self):
        return self._D

    @property
    def D(self):
        return self._D

    @property
    def requires_grad(self):
        return self._F.requires_grad

    def requires_grad_(self, requires_grad: bool = True):
        self._F.requires_grad_(requires_grad)

    def float(self):
        self._F = self._F.float()
        return self

    def double(self):
        self._F = self._F.double()
        return self

    def __len__(self):
        return len(self._F)

    def size(self):
        return self._F.size()

    @property
    def shape(self):
        return self._F.shape

    @property
    def device(self):
        return self._F.device

    @property
    def dtype(self):
        return self._F.dtype

    def detach(self):
        self._F = self._F.detach()
        return self

    def get_device(self):
        return self._F.get_device()

    # -- arithmetic ------------------------------------------------------------------
    def _is_same_key(self, other):
        assert isinstance(other, self.__class__)
        assert self._manager == other._manager, COORDINATE_MANAGER_DIFFERENT_ERROR
        assert self.coordinate_map_key == other.coordinate_map_key, COORDINATE_KEY_DIFFERENT_ERROR

    def __iadd__(self, other):
        self._is_same_key(other)
        self._F += other.F
        return self

    def __isub__(self, other):
        self._is_same_key(other)
        self._F -= other.F
        return self

    def __imul__(self, other):
        self._is_same_key(other)
        self._F *= other.F
        return self

    def __idiv__(self, other):
        self._is_same_key(other)
        self._F /= other.F
        return self

    def _binary_functor(self, other, binary_fn):
        assert isinstance(other, (self.__class__, torch.Tensor))
        if not isinstance(other, self.__class__):           # plain torch.Tensor / broadcastable
            return self.__class__(binary_fn(self._F, other), coordinate_map_key=self.coordinate_map_key,
                                  coordinate_manager=self._manager)
        assert self._manager == other._manager, COORDINATE_MANAGER_DIFFERENT_ERROR
        if self.coordinate_map_key == other.coordinate_map_key:
            return self.__class__(binary_fn(self._F, other.F), coordinate_map_key=self.coordinate_map_key,
                                  coordinate_manager=self._manager)
        # different maps: operate on the union, the missing side counts as 0
        out_key = CoordinateMapKey(self.coordinate_map_key.get_coordinate_size())
        maps = self.coordinate_manager.union_map([self.coordinate_map_key, other.coordinate_map_key], out_key)
        n_out = self.coordinate_manager.size(out_key)
        out_F = torch.zeros((n_out, self._F.size(1)), dtype=self.dtype, device=self.device)
        out_F[maps[0][1]] = self._F[maps[0][0]]
        out_F[maps[1][1]] = binary_fn(out_F[maps[1][1]], other._F[maps[1][0]])
        return self.__class__(out_F, coordinate_map_key=out_key, coordinate_manager=self._manager)

    def __add__(self, other):
        return self._binary_functor(other, lambda x, y: x + y)

    def __sub__(self, other):
        return self._binary_functor(other, lambda x, y: x - y)

    def __mul__(self, other):
        return self._binary_functor(other, lambda x, y: x * y)

    def __truediv__(self, other):
        return self._binary_functor(other, lambda x, y: x / y)

    def __power__(self, power):
        return self.__class__(self._F ** power, coordinate_map_key=self.coordinate_map_key,
                              coordinate_manager=self._manager)
