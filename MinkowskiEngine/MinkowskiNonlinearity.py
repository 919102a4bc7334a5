"""Element-wise layers applied to the feature matrix (reference:
MinkowskiNonlinearity.py:34-200): every `torch.nn` activation `X` is exposed as
`MinkowskiX` with the same constructor arguments."""
import torch
import torch.nn

from .MinkowskiCommon import MinkowskiModuleBase
from .MinkowskiSparseTensor import SparseTensor


class MinkowskiNonlinearityBase(MinkowskiModuleBase):
    MODULE = None

    def __init__(self, *args, **kwargs):
        super().__init__()
        self.module = self.MODULE(*args, **kwargs)

    def forward(self, input):
        return SparseTensor(self.module(input.F), coordinate_map_key=input.coordinate_map_key,
                            coordinate_manager=input.coordinate_manager)

    def __repr__(self):
        return self.__class__.__name__ + "()"


_WRAPPED = ["ELU", "Hardshrink", "Hardsigmoid", "Hardtanh", "Hardswish", "LeakyReLU", "LogSigmoid", "PReLU", "ReLU",
            "ReLU6", "RReLU", "SELU", "CELU", "GELU", "Sigmoid", "SiLU", "Softplus", "Softshrink", "Softsign",
            "Tanh", "Tanhshrink", "Threshold", "Softmin", "Softmax", "LogSoftmax", "AdaptiveLogSoftmaxWithLoss",
            "Dropout", "AlphaDropout"]
for _name in _WRAPPED:
    globals()["Minkowski" + _name] = type("Minkowski" + _name, (MinkowskiNonlinearityBase,),
                                          {"MODULE": getattr(torch.nn, _name), "__module__": __name__})
del _name


class MinkowskiSinusoidal(MinkowskiModuleBase):
    """sin(x @ kernel + bias) positional features."""

    def __init__(self, in_channel, out_channel):
        MinkowskiModuleBase.__init__(self)
        self.in_channel = in_channel
        self.out_channel = out_channel
        self.kernel = torch.nn.Parameter(torch.rand(in_channel, out_channel))
        self.bias = torch.nn.Parameter(torch.rand(1, out_channel))
        self.coef = torch.nn.Parameter(torch.rand(1, out_channel))

    def forward(self, input):
        out_F = torch.sin(input.F.mm(self.kernel) + self.bias) * self.coef
        return SparseTensor(out_F, coordinate_map_key=input.coordinate_map_key,
                            coordinate_manager=input.coordinate_manager)
