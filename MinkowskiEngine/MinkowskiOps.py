"""Feature-level helpers (reference: MinkowskiOps.py:40-497): linear layer,
cat / sum / mean / var over tensors on the same map, dense <-> sparse."""
from typing import Union

import torch
from torch.nn import Module

from MinkowskiEngineBackend._C import CoordinateMapKey

from .MinkowskiCommon import MinkowskiModuleBase
from .MinkowskiSparseTensor import SparseTensor
from .MinkowskiTensor import COORDINATE_KEY_DIFFERENT_ERROR, COORDINATE_MANAGER_DIFFERENT_ERROR


class MinkowskiLinear(Module):
    def __init__(self, in_features, out_features, bias=True):
        super().__init__()
        self.linear = torch.nn.Linear(in_features, out_features, bias=bias)

    def forward(self, input: SparseTensor):
        return SparseTensor(self.linear(input.F), coordinate_map_key=input.coordinate_map_key,
                            coordinate_manager=input.coordinate_manager)

    def __repr__(self):
        return self.__class__.__name__ + (f"(in_features={self.linear.in_features}, "
                                          f"out_features={self.linear.out_features}, "
                                          f"bias={self.linear.bias is not None})")


def _tuple_operator(*sparse_tensors, operator):
    if len(sparse_tensors) == 1:
        assert isinstance(sparse_tensors[0], (tuple, list))
        sparse_tensors = sparse_tensors[0]
    assert len(sparse_tensors) > 1, "Invalid number of inputs. The input must be at least two len(sparse_tensors) > 1"
    first = sparse_tensors[0]
    assert isinstance(first, SparseTensor), "Invalid data type. The input must be a list of sparse tensors."
    for s in sparse_tensors:
        assert isinstance(s, SparseTensor), "Inputs must be SparseTensors."
        assert first.device == s.device, f"Device must be the same. {first.device} != {s.device}"
        assert first.coordinate_manager == s.coordinate_manager, COORDINATE_MANAGER_DIFFERENT_ERROR
        assert first.coordinate_map_key == s.coordinate_map_key, \
            COORDINATE_KEY_DIFFERENT_ERROR + str(first.coordinate_map_key) + " != " + str(s.coordinate_map_key)
    return SparseTensor(operator([s.F for s in sparse_tensors]), coordinate_map_key=first.coordinate_map_key,
                        coordinate_manager=first.coordinate_manager)


def cat(*sparse_tensors):
    """Concatenate features of sparse tensors that share a coordinate map."""
    return _tuple_operator(*sparse_tensors, operator=lambda xs: torch.cat(xs, dim=-1))


def _sum(*sparse_tensors):
    return _tuple_operator(*sparse_tensors, operator=lambda xs: torch.stack(xs).sum(0))


def mean(*sparse_tensors):
    return _tuple_operator(*sparse_tensors, operator=lambda xs: torch.stack(xs).mean(0))


def var(*sparse_tensors):
    return _tuple_operator(*sparse_tensors, operator=lambda xs: torch.stack(xs).var(0, unbiased=False))


def dense_coordinates(shape: Union[list, torch.Size]):
    """Coordinates [B * prod(spatial), D+1] of every cell of a dense [B, C, *spatial] tensor."""
    spatial = list(shape[2:])
    B = shape[0]
    grids = torch.meshgrid(*[torch.arange(s, dtype=torch.int32) for s in [B] + spatial], indexing="ij")
    return torch.stack([g.reshape(-1) for g in grids], dim=1).contiguous()


def to_sparse(x: torch.Tensor, format: str = None, coordinates=None, device=None):
    """Dense [B, C, *spatial] -> SparseTensor over the non-zero cells."""
    assert x.ndim > 2, "Input has 0 spatial dimension."
    assert isinstance(x, torch.Tensor)
    if format is None:
        format = ["B", "C"] + ["X"] * (x.ndim - 2)
    assert x.ndim == len(format), f"Invalid format: {format}. len(format) != x.ndim"
    assert "B" in format and "B" == format[0] and format.count("B") == 1, "the input must have the batch axis first"
    assert "C" in format and format.count("C") == 1, "The format must contain 'C' for channel"
    if device is None:
        device = x.device
    ch_dim = format.index("C")
    reduced = torch.abs(x).sum(ch_dim)
    nz = torch.where(reduced != 0)
    perm = [i for i in range(x.ndim) if i != ch_dim] + [ch_dim]
    feats = x.permute(*perm)[nz]
    coords = torch.stack(nz, dim=1).int()
    return SparseTensor(features=feats.to(device), coordinates=coords.to(device), device=device)


def to_sparse_all(dense_tensor: torch.Tensor, coordinates: torch.Tensor = None):
    """Dense [B, C, *spatial] -> SparseTensor over *all* cells."""
    assert dense_tensor.ndim > 2
    if coordinates is None:
        coordinates = dense_coordinates(dense_tensor.shape)
    C = dense_tensor.shape[1]
    perm = [0] + list(range(2, dense_tensor.ndim)) + [1]
    feats = dense_tensor.permute(*perm).reshape(-1, C)
    return SparseTensor(feats, coordinates.to(feats.device), device=dense_tensor.device)


class MinkowskiToSparseTensor(MinkowskiModuleBase):
    def __init__(self, remove_zeros=True, coordinates: torch.Tensor = None):
        MinkowskiModuleBase.__init__(self)
        assert remove_zeros or coordinates is None or True
        self.remove_zeros = remove_zeros
        self.coordinates = coordinates

    def forward(self, input: torch.Tensor):
        if isinstance(input, torch.Tensor):
            return to_sparse(input) if self.remove_zeros else to_sparse_all(input, self.coordinates)
        raise ValueError("Unsupported type. Only torch.Tensor is supported")

    def __repr__(self):
        return self.__class__.__name__ + "()"


class MinkowskiToDenseTensor(MinkowskiModuleBase):
    def __init__(self, shape: torch.Size = None):
        MinkowskiModuleBase.__init__(self)
        self.shape = shape

    def forward(self, input: SparseTensor):
        dense, _, _ = input.dense(shape=self.shape)
        return dense

    def __repr__(self):
        return self.__class__.__name__ + "()"


class MinkowskiToFeature(MinkowskiModuleBase):
    def forward(self, x: SparseTensor):
        assert isinstance(x, SparseTensor), "Invalid input type for MinkowskiToFeature"
        return x.F


class MinkowskiStackCat(torch.nn.Sequential):
    def forward(self, x):
        return cat([module(x) for module in self])


class MinkowskiStackSum(torch.nn.Sequential):
    def forward(self, x):
        return _sum([module(x) for module in self])


class MinkowskiStackMean(torch.nn.Sequential):
    def forward(self, x):
        return mean([module(x) for module in self])


class MinkowskiStackVar(torch.nn.Sequential):
    def forward(self, x):
        return var([module(x) for module in self])
