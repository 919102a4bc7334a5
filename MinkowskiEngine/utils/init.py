"""Kaiming initialisation for kernels laid out [K, Cin, Cout] (reference: utils/init.py)."""
import math

import torch


def _calculate_fan_in_and_fan_out(tensor):
    if tensor.dim() < 2:
        raise ValueError("Fan in and fan out can not be computed for tensor with fewer than 2 dimensions")
    if tensor.dim() == 2:   # linear / kernel volume 1: [in, out] seen as torch's [out, in] convention
        return tensor.size(1), tensor.size(0)
    k = tensor.size(0)      # receptive field = kernel volume
    return tensor.size(1) * k, tensor.size(2) * k


def _calculate_correct_fan(tensor, mode):
    mode = mode.lower()
    if mode not in ("fan_in", "fan_out"):
        raise ValueError(f"Mode {mode} not supported, please use one of ['fan_in', 'fan_out']")
    fan_in, fan_out = _calculate_fan_in_and_fan_out(tensor)
    return fan_in if mode == "fan_in" else fan_out


def kaiming_normal_(tensor, a=0, mode="fan_in", nonlinearity="leaky_relu"):
    fan = _calculate_correct_fan(tensor, mode)
    std = torch.nn.init.calculate_gain(nonlinearity, a) / math.sqrt(fan)
    with torch.no_grad():
        return tensor.normal_(0, std)
