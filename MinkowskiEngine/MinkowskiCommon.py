"""Shared helpers of the ME Python API (mirrors the public names of the
reference's MinkowskiCommon.py:25-120)."""
from collections.abc import Sequence
from typing import Union

import numpy as np
import torch
from torch.nn import Module

import MinkowskiEngineBackend._C as MEB

StrideType = Union[int, Sequence, np.ndarray, torch.IntTensor]


def convert_to_int_list(arg, dimension: int):
    """int | sequence | array | tensor -> list of `dimension` ints."""
    if isinstance(arg, list):
        assert len(arg) == dimension, f"expected {dimension} values, got {arg}"
        return arg
    if isinstance(arg, (Sequence, np.ndarray, torch.Tensor)):
        out = [int(v) for v in arg]
        assert len(out) == dimension, f"expected {dimension} values, got {arg}"
        return out
    if np.isscalar(arg):
        return [int(arg)] * dimension
    raise ValueError("Input must be a scalar or a sequence")


def convert_to_int_tensor(arg, dimension: int):
    if isinstance(arg, torch.IntTensor):
        assert arg.numel() == dimension
        return arg
    return torch.IntTensor(convert_to_int_list(arg, dimension))


def prep_args(tensor_stride, stride, kernel_size, dilation, region_type, D=-1):
    assert D > 0, f"dimension must be a positive integer, {D}"
    return (convert_to_int_tensor(tensor_stride, D), convert_to_int_tensor(stride, D),
            convert_to_int_tensor(kernel_size, D), convert_to_int_tensor(dilation, D), int(region_type))


def get_postfix(tensor: torch.Tensor) -> str:
    return "GPU" if tensor.is_cuda else "CPU"


class MinkowskiModuleBase(Module):
    pass


def get_minkowski_function(name: str, variable: torch.Tensor):
    """Resolve `name + {GPU|CPU}` in the backend (reference: MinkowskiCommon.py:101-120).
    This engine exports only *GPU functions; CPU tensors raise."""
    fn_name = name + get_postfix(variable)
    fn = getattr(MEB, fn_name, None)
    if fn is None:
        if variable.is_cuda:
            raise ValueError(f"Function {fn_name} not available.")
        raise ValueError(f"Function {fn_name} not available: this engine is CUDA-only (sm_100a); "
                         "move the SparseTensor to a CUDA device.")
    return fn
