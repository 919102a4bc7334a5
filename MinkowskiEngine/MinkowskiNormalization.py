"""Normalisation layers on sparse tensors (reference: MinkowskiNormalization.py).

Batch norms wrap torch's BatchNorm1d / SyncBatchNorm on the feature matrix
(:55-195).  The instance-norm family (:198-622, including the fork's grouped /
adaptive variants) is built from the engine's global average pooling and
broadcast kernels: per batch item  y = (x - mean_b) / sqrt(var_b + eps).
Parameter names and shapes match the reference so checkpoints load.
"""
import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.nn import Module, init

from MinkowskiEngineBackend._C import BroadcastMode, PoolingMode

from .MinkowskiBroadcast import (MinkowskiBroadcastAddition, MinkowskiBroadcastFunction,
                                 MinkowskiBroadcastMultiplication)
from .MinkowskiCommon import MinkowskiModuleBase
from .MinkowskiPooling import MinkowskiGlobalAvgPooling, MinkowskiGlobalPoolingFunction
from .MinkowskiSparseTensor import SparseTensor


def _like(x: SparseTensor, feats: torch.Tensor) -> SparseTensor:
    return SparseTensor(feats, coordinate_map_key=x.coordinate_map_key, coordinate_manager=x.coordinate_manager)


# ---------------------------------------------------------------------------
# fused paths on the engine (CUDA features): one statistics pass + one affine pass
# (mink_octtree_stablediffusion_b200/backend_norm.py, csrc/norm.cu).  CPU tensors keep the
# composed form below (torch batch norm; global pooling + broadcast functions).
# ---------------------------------------------------------------------------
def _acc(t: torch.Tensor) -> torch.dtype:
    return torch.float64 if t.dtype == torch.float64 else torch.float32


def _bn_coef(eps):
    def coef(mean, var, weight, bias):
        inv = torch.rsqrt(var + eps)
        a = inv if weight is None else inv * weight.to(inv.dtype)
        c = -mean * a
        if bias is not None:
            c = c + bias.to(inv.dtype)
        return a, c
    return coef


def _fused_batch_norm(bn: torch.nn.modules.batchnorm._BatchNorm, x: torch.Tensor, sync_group=False):
    """Training-mode batch norm of the feature matrix (statistics over all rows; for
    SyncBatchNorm over the rows of all ranks), running statistics updated like torch's
    (`momentum`, unbiased running variance).  Returns None when the case is not served
    (eval mode, no rows, parameters on another device) and the caller uses torch's own."""
    from mink_octtree_stablediffusion_b200.backend_norm import batch_norm_train
    use_batch_stats = bn.training or not bn.track_running_stats or bn.running_mean is None
    if not use_batch_stats or x.shape[0] == 0 or not x.is_floating_point():
        return None
    acc = _acc(x)
    tensors = [t for t in (bn.weight, bn.bias, bn.running_mean, bn.running_var) if t is not None]
    if any(not t.is_floating_point() or t.device != x.device for t in tensors):
        return None
    reduce_fn = None
    if sync_group is not False:
        import torch.distributed as dist
        group = sync_group
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            def reduce_fn(flat):
                dist.all_reduce(flat, group=group)
    track = bn.training and bn.track_running_stats and bn.running_mean is not None
    momentum = None
    if track:
        bn.num_batches_tracked.add_(1)
        # momentum=None is torch's cumulative average (needs the step count on the host)
        momentum = bn.momentum if bn.momentum is not None else 1.0 / float(bn.num_batches_tracked)
    return batch_norm_train(x, bn.weight, bias_or_none(bn), bn.running_mean if track else None,
                            bn.running_var if track else None, momentum, bn.eps, reduce_fn)


def bias_or_none(bn):
    return bn.bias if bn.affine else None


def _segments(x: SparseTensor):
    """(seg [N] int32 batch row of every voxel, n_batch, counts [B], first row of every batch item),
    cached on the manager beside the coordinate map."""
    mgr = x.coordinate_manager._manager
    cache = mgr.__dict__.setdefault("_norm_segments", {})
    k = x.coordinate_map_key._tuple()
    hit = cache.get(k)
    if hit is None:
        rows, nb = mgr._batch_of_row(x.coordinate_map_key)
        r = rows.long()
        counts = torch.bincount(r.clamp(min=0), minlength=nb).to(torch.float32)
        first = torch.full((nb,), rows.numel(), dtype=torch.int64, device=rows.device)
        first.scatter_reduce_(0, r.clamp(min=0), torch.arange(rows.numel(), device=rows.device), reduce="amin")
        hit = cache[k] = (rows, nb, counts, first.clamp(max=max(rows.numel() - 1, 0)))
    return hit


def _fused_instance_norm(x: SparseTensor, coef_fn, params):
    from mink_octtree_stablediffusion_b200.backend_norm import segment_norm
    seg, nb, counts, first = _segments(x)
    F_ = x.F
    shift = F_.detach()[first].to(_acc(F_)).contiguous()
    return segment_norm(F_, seg, nb, counts.to(_acc(F_)), coef_fn, params, shift=shift)


def _fused_group_norm(x: SparseTensor, weight, bias, eps, gs: int, gw: int):
    """Instance / group norm with its coefficient algebra in one launch each way (csrc/norm.cu
    `mk_gn_coef_fwd/bwd`): statistics over a batch item's rows and `gs` adjacent channels, affine
    parameters shared by `gw` adjacent channels."""
    from mink_octtree_stablediffusion_b200.backend_norm import group_instance_norm
    seg, nb, counts, first = _segments(x)
    F_ = x.F
    acc = _acc(F_)
    shift = F_.detach()[first].to(acc).contiguous()
    return group_instance_norm(F_, seg, nb, counts if counts.dtype == acc else counts.to(acc), shift, weight, bias,
                               eps, gs, gw)


def _fold_group(stat, group):
    if group == 1:
        return stat
    c = stat.shape[-1]
    g = stat.view(-1, c // group, group).mean(dim=-1)
    return torch.repeat_interleave(g, group, dim=-1)


def _stable_coef(eps, fold):
    """Coefficients of the fork's 'stable' norms from the per-channel statistics: the data are
    centred with the FOLDED mean mu_g, so the per-channel second moment about it is
    var_c + (mean_c - mu_g)^2, which is then folded as well (MinkowskiNormalization.py:347-402)."""
    def coef(mean, var, weight, bias, *ada):
        mu = fold(mean)
        v = fold(var + (mean - mu) ** 2)
        inv = torch.rsqrt(v + eps)
        a = inv * weight.to(inv.dtype)
        c = bias.to(inv.dtype) - mu * a
        if ada:                      # AdaIN: (y w + b) (1 + scale[batch]) + shift[batch]
            scale, shift = ada
            a = a * (1 + scale.to(inv.dtype))
            c = c * (1 + scale.to(inv.dtype)) + shift.to(inv.dtype)
        return a, c
    return coef


class MinkowskiBatchNorm(Module):
    """torch.nn.BatchNorm1d over the feature rows of a sparse tensor."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True):
        super().__init__()
        self.bn = torch.nn.BatchNorm1d(num_features, eps=eps, momentum=momentum, affine=affine,
                                       track_running_stats=track_running_stats)

    def forward(self, input):
        x = input.F
        y = _fused_batch_norm(self.bn, x) if x.is_cuda else None
        return _like(input, self.bn(x) if y is None else y)

    def __repr__(self):
        bn = self.bn
        return self.__class__.__name__ + (f"({bn.num_features}, eps={bn.eps}, momentum={bn.momentum}, "
                                          f"affine={bn.affine}, track_running_stats={bn.track_running_stats})")


class MinkowskiSyncBatchNorm(MinkowskiBatchNorm):
    """torch.nn.SyncBatchNorm over the feature rows (statistics all-gathered over
    the process group, NCCL under DDP)."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True,
                 process_group=None):
        Module.__init__(self)
        self.bn = torch.nn.SyncBatchNorm(num_features, eps=eps, momentum=momentum, affine=affine,
                                         track_running_stats=track_running_stats, process_group=process_group)

    def forward(self, input):
        x = input.F
        # NB `convert_sync_batchnorm` below re-attaches the ORIGINAL BatchNorm1d as `.bn` when it recurses into the
        # converted module's children - the reference does exactly that (MinkowskiNormalization.py:188-192), so
        # a converted network normalises with local statistics there and here; a directly constructed
        # MinkowskiSyncBatchNorm synchronises.
        group = self.bn.process_group if isinstance(self.bn, torch.nn.SyncBatchNorm) else False
        y = _fused_batch_norm(self.bn, x, sync_group=group) if x.is_cuda else None
        return _like(input, self.bn(x) if y is None else y)

    @classmethod
    def convert_sync_batchnorm(cls, module, process_group=None):
        """Recursively replace every MinkowskiBatchNorm by MinkowskiSyncBatchNorm,
        keeping affine parameters and running statistics."""
        out = module
        if isinstance(module, MinkowskiBatchNorm) and not isinstance(module, MinkowskiSyncBatchNorm):
            src = module.bn
            out = MinkowskiSyncBatchNorm(src.num_features, src.eps, src.momentum, src.affine,
                                         src.track_running_stats, process_group)
            if src.affine:
                with torch.no_grad():
                    out.bn.weight = src.weight
                    out.bn.bias = src.bias
            out.bn.running_mean = src.running_mean
            out.bn.running_var = src.running_var
            out.bn.num_batches_tracked = src.num_batches_tracked
            if hasattr(src, "qconfig"):
                out.bn.qconfig = src.qconfig
        for name, child in module.named_children():
            out.add_module(name, cls.convert_sync_batchnorm(child, process_group))
        del module
        return out


def _instance_standardise(x: SparseTensor, eps: float, fold=None):
    """(x - mean_b) / sqrt(var_b + eps) per batch item b.  `fold` optionally
    merges the per-channel statistics across channel groups ([B, C] -> [B, C])."""
    mgr, key = x.coordinate_manager, x.coordinate_map_key
    avg = PoolingMode.GLOBAL_AVG_POOLING_PYTORCH_INDEX
    mean = MinkowskiGlobalPoolingFunction.apply(x.F, avg, key, None, mgr)
    okey = mgr.origin()
    if fold is not None:
        mean = fold(mean)
    centered = MinkowskiBroadcastFunction.apply(x.F, -mean, BroadcastMode.ELEMENTWISE_ADDITON, key, okey, mgr)
    var = MinkowskiGlobalPoolingFunction.apply(centered * centered, avg, key, None, mgr)
    if fold is not None:
        var = fold(var)
    inv_std = 1.0 / (var + eps).sqrt()
    return MinkowskiBroadcastFunction.apply(centered, inv_std, BroadcastMode.ELEMENTWISE_MULTIPLICATION, key, okey,
                                            mgr)


class MinkowskiInstanceNorm(MinkowskiModuleBase):
    """Instance normalisation with per-channel affine (weight/bias [1, C])."""

    def __init__(self, num_features):
        Module.__init__(self)
        self.num_features = num_features
        self.eps = 1e-6
        self.weight = nn.Parameter(torch.ones(1, num_features))
        self.bias = nn.Parameter(torch.zeros(1, num_features))
        self.reset_parameters()

    def __repr__(self):
        return self.__class__.__name__ + f"(nchannels={self.num_features})"

    def reset_parameters(self):
        self.weight.data.fill_(1)
        self.bias.data.zero_()

    def forward(self, input: SparseTensor):
        assert isinstance(input, SparseTensor)
        if input.F.is_cuda:
            return _like(input, _fused_group_norm(input, self.weight, self.bias, self.eps, 1, 1))
        y = _instance_standardise(input, self.eps)
        return _like(input, y * self.weight + self.bias)


class MinkowskiStableInstanceNorm(MinkowskiModuleBase):
    """Fork variant (:317-402): statistics are averaged over `group` adjacent
    channels and the affine parameters are shared inside a group
    (weight/bias [1, C // group])."""

    def __init__(self, num_features, group=1):
        Module.__init__(self)
        self.num_features = num_features
        if group > num_features:
            print("group should be less than num_features, set group=num_features")
            group = num_features
        self.group = group
        assert self.num_features % self.group == 0
        self.eps = 1e-6
        self.weight = nn.Parameter(torch.ones(1, num_features // self.group))
        self.bias = nn.Parameter(torch.zeros(1, num_features // self.group))
        # kept as submodules for state_dict / repr parity with the reference
        self.mean_in = MinkowskiGlobalAvgPooling()
        self.glob_sum = MinkowskiBroadcastAddition()
        self.glob_sum2 = MinkowskiBroadcastAddition()
        self.glob_mean = MinkowskiGlobalAvgPooling()
        self.glob_times = MinkowskiBroadcastMultiplication()
        self.reset_parameters()

    def __repr__(self):
        return self.__class__.__name__ + f"(nchannels={self.num_features})"

    def reset_parameters(self):
        self.weight.data.fill_(1)
        self.bias.data.zero_()

    def _fold(self, stat):
        if self.group == 1:
            return stat
        g = stat.view(-1, self.num_features // self.group, self.group).mean(dim=-1)
        return torch.repeat_interleave(g, self.group, dim=-1)

    def forward(self, x):
        if x.F.is_cuda:
            return _like(x, _fused_group_norm(x, self.weight, self.bias, self.eps, self.group, self.group))
        y = _instance_standardise(x, self.eps, self._fold)
        scale = torch.repeat_interleave(self.weight, self.group, dim=-1)
        shift = torch.repeat_interleave(self.bias, self.group, dim=-1)
        return _like(x, y * scale + shift)


class MinkowskiStableGroupNorm(MinkowskiModuleBase):
    """Fork variant (:405-478): one mean / variance per batch item across all
    channels (layer-norm-like), per-channel affine."""

    def __init__(self, num_features, group=None):
        Module.__init__(self)
        self.num_features = num_features
        self.eps = 1e-6
        self.weight = nn.Parameter(torch.ones(1, num_features))
        self.bias = nn.Parameter(torch.zeros(1, num_features))
        self.mean_in = MinkowskiGlobalAvgPooling()
        self.glob_sum = MinkowskiBroadcastAddition()
        self.glob_sum2 = MinkowskiBroadcastAddition()
        self.glob_mean = MinkowskiGlobalAvgPooling()
        self.glob_times = MinkowskiBroadcastMultiplication()
        self.reset_parameters()

    def __repr__(self):
        return self.__class__.__name__ + f"(nchannels={self.num_features})"

    def reset_parameters(self):
        self.weight.data.fill_(1)
        self.bias.data.zero_()

    def forward(self, x):
        fold = lambda s: s.mean(dim=1, keepdim=True).repeat(1, self.num_features)  # noqa: E731
        if x.F.is_cuda:
            return _like(x, _fused_group_norm(x, self.weight, self.bias, self.eps, self.num_features, 1))
        y = _instance_standardise(x, self.eps, fold)
        return _like(x, y * self.weight + self.bias)


class HjmInstanceNorm(MinkowskiBatchNorm):
    """Fork variant (:520-543): BatchNorm1d applied to each batch item separately."""

    def forward(self, input):
        feats = input.F
        output = torch.zeros_like(feats)
        for rows in input.decomposition_permutations:
            if rows.numel() > 0:
                output[rows] = output[rows] + self.bn(feats[rows])
        return _like(input, output)


class AdaStableInstanceNorm(MinkowskiModuleBase):
    """Fork variant (:546-622): instance norm followed by an embedding-conditioned
    scale/shift (AdaIN): (y*w + b) * (1 + scale[batch]) + shift[batch]."""

    def __init__(self, num_features, embedding_dim):
        Module.__init__(self)
        self.num_features = num_features
        self.eps = 1e-6
        self.weight = nn.Parameter(torch.ones(1, num_features))
        self.bias = nn.Parameter(torch.zeros(1, num_features))
        self.mean_in = MinkowskiGlobalAvgPooling()
        self.glob_sum = MinkowskiBroadcastAddition()
        self.glob_sum2 = MinkowskiBroadcastAddition()
        self.glob_mean = MinkowskiGlobalAvgPooling()
        self.glob_times = MinkowskiBroadcastMultiplication()
        self.linear = nn.Linear(embedding_dim, num_features * 2)
        self.reset_parameters()

    def __repr__(self):
        return self.__class__.__name__ + f"(nchannels={self.num_features})"

    def reset_parameters(self):
        self.weight.data.fill_(1)
        self.bias.data.zero_()
        nn.init.normal_(self.linear.weight, 0, 0.01)
        nn.init.constant_(self.linear.bias, 0)

    def forward(self, x, emb):
        scale, shift = self.linear(emb).chunk(2, dim=1)
        if x.F.is_cuda and scale.shape[0] == _segments(x)[1]:
            # scale / shift rows are indexed by batch INDEX; the segments are the rows of the origin map,
            # which are ordered by batch index (row b == batch b when the indices are 0..B-1)
            return _like(x, _fused_instance_norm(x, _stable_coef(self.eps, lambda s: s),
                                                 (self.weight, self.bias, scale, shift)))
        y = _instance_standardise(x, self.eps)
        b = x.C[:, 0].long()
        return _like(x, (y * self.weight + self.bias) * (1 + scale[b]) + shift[b])


class HjmGroupNorm(Module):
    """Fork variant (:624-711): group norm on a dense tensor with one affine
    scalar per group."""

    def __init__(self, num_groups: int, num_channels: int, eps: float = 1e-5, affine: bool = True,
                 device=None, dtype=None) -> None:
        factory_kwargs = {"device": device, "dtype": dtype}
        super().__init__()
        if num_channels % num_groups != 0:
            raise ValueError("num_channels must be divisible by num_groups")
        self.num_groups = num_groups
        self.num_channels = num_channels
        self.eps = eps
        self.affine = affine
        if self.affine:
            self.weight = nn.Parameter(torch.empty(num_groups, **factory_kwargs))
            self.bias = nn.Parameter(torch.empty(num_groups, **factory_kwargs))
        else:
            self.register_parameter("weight", None)
            self.register_parameter("bias", None)
        self.reset_parameters()

    def reset_parameters(self) -> None:
        if self.affine:
            init.ones_(self.weight)
            init.zeros_(self.bias)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        rep = self.num_channels // self.num_groups
        weight = torch.repeat_interleave(self.weight, rep, dim=-1) if self.affine else None
        bias = torch.repeat_interleave(self.bias, rep, dim=-1) if self.affine else None
        return F.group_norm(input, self.num_groups, weight, bias, self.eps)

    def extra_repr(self) -> str:
        return f"{self.num_groups}, {self.num_channels}, eps={self.eps}, affine={self.affine}"
