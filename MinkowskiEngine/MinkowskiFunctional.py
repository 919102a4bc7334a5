"""`torch.nn.functional` on sparse tensors (reference: MinkowskiFunctional.py:30-232):
activations return a SparseTensor on the same map, losses act on `.F`."""
import torch.nn.functional as F

from .MinkowskiSparseTensor import SparseTensor


def _wrap_tensor(input, feats):
    return SparseTensor(feats, coordinate_map_key=input.coordinate_map_key,
                        coordinate_manager=input.coordinate_manager)


def _unary(name):
    fn = getattr(F, name)

    def op(input, *args, **kwargs):
        return _wrap_tensor(input, fn(input.F, *args, **kwargs))
    op.__name__ = name
    return op


def _loss(name):
    fn = getattr(F, name)

    def op(input, target, *args, **kwargs):
        return fn(input.F, target, *args, **kwargs)
    op.__name__ = name
    return op


for _n in ["threshold", "relu", "hardtanh", "hardswish", "relu6", "elu", "selu", "celu", "leaky_relu", "prelu",
           "rrelu", "glu", "gelu", "logsigmoid", "hardshrink", "tanhshrink", "softsign", "softplus", "softmin",
           "softmax", "softshrink", "gumbel_softmax", "log_softmax", "tanh", "sigmoid", "hardsigmoid", "silu",
           "batch_norm", "normalize", "linear", "dropout", "alpha_dropout"]:
    globals()[_n] = _unary(_n)
for _n in ["binary_cross_entropy", "binary_cross_entropy_with_logits", "poisson_nll_loss", "cross_entropy",
           "hinge_embedding_loss", "kl_div", "l1_loss", "mse_loss", "multilabel_margin_loss",
           "multilabel_soft_margin_loss", "multi_margin_loss", "nll_loss", "smooth_l1_loss", "soft_margin_loss"]:
    globals()[_n] = _loss(_n)
del _n
