"""Synchronised BatchNorm: batch statistics over ALL replicas of a data-parallel job.

Reference: model/cv/batchnorm_utils.py:156-266 (``SynchronizedBatchNorm{1,2,3}d`` for ``nn.DataParallel`` with a
master / slave pipe between replica threads, plus ``convert_model``) used by the segmentation models.  Here replicas are
processes (one per GPU) of a ``torch.distributed`` group: each forward all-reduces per-channel ``(sum, sum of squares,
count)``, each backward all-reduces ``(sum dy, sum dy * xhat)``, so the normalisation and its gradient equal BatchNorm on
the concatenated global batch.  Works on any backend (NCCL over NVLink on the GPUs, gloo in the CPU tests); with no
process group (or world size 1) it is plain BatchNorm.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn


def _group_ok(group) -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


class _SyncBNFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps, group):
        C = x.shape[1]
        dims = [0] + list(range(2, x.dim()))
        xf = x if x.dtype == torch.float64 else x.float()      # statistics in fp32 (fp64 inputs stay fp64)
        n_local = x.numel() // C
        stats = torch.cat([xf.sum(dims), (xf * xf).sum(dims), torch.tensor([float(n_local)], device=x.device, dtype=xf.dtype)])
        if _group_ok(group):
            dist.all_reduce(stats, group=group)
        n = stats[-1]
        mean = stats[:C] / n
        var = (stats[C:2 * C] / n - mean * mean).clamp_min(0.0)
        rstd = torch.rsqrt(var + eps)
        shape = [1, C] + [1] * (x.dim() - 2)
        xhat = (xf - mean.view(shape)) * rstd.view(shape)
        y = xhat
        if weight is not None:
            y = y * weight.to(xf.dtype).view(shape) + bias.to(xf.dtype).view(shape)
        ctx.save_for_backward(xhat, rstd, weight if weight is not None else torch.ones(C, device=x.device))
        ctx.group, ctx.n, ctx.dims, ctx.shape, ctx.affine = group, n, dims, shape, weight is not None
        ctx.mark_non_differentiable(mean, var)
        return y.to(x.dtype), mean, var, n

    @staticmethod
    def backward(ctx, dy, _dm, _dv, _dn):
        xhat, rstd, weight = ctx.saved_tensors
        C = xhat.shape[1]
        dyf = dy.to(xhat.dtype)
        sums = torch.cat([dyf.sum(ctx.dims), (dyf * xhat).sum(ctx.dims)])
        dbeta_local, dgamma_local = sums[:C].clone(), sums[C:].clone()
        if _group_ok(ctx.group):
            dist.all_reduce(sums, group=ctx.group)
        sd, sdx = sums[:C] / ctx.n, sums[C:] / ctx.n
        g = (weight.to(xhat.dtype) * rstd).view(ctx.shape)
        dx = g * (dyf - sd.view(ctx.shape) - xhat * sdx.view(ctx.shape))
        if ctx.affine:
            return dx.to(dy.dtype), dgamma_local.to(weight.dtype), dbeta_local.to(weight.dtype), None, None
        return dx.to(dy.dtype), None, None, None, None


class SynchronizedBatchNorm(nn.modules.batchnorm._BatchNorm):
    """Drop-in for ``nn.BatchNorm{1,2,3}d`` (same parameters, buffers and ``state_dict`` keys)."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True, process_group=None):
        super().__init__(num_features, eps, momentum, affine, track_running_stats)
        self.process_group = process_group

    def _check_input_dim(self, x):
        if x.dim() < 2:
            raise ValueError(f"expected at least 2D input (got {x.dim()}D)")

    def forward(self, x):
        self._check_input_dim(x)
        if not self.training and self.track_running_stats:
            return nn.functional.batch_norm(x, self.running_mean, self.running_var, self.weight, self.bias, False, 0.0, self.eps)
        y, mean, var, n = _SyncBNFunction.apply(x, self.weight, self.bias, self.eps, self.process_group)
        if self.training and self.track_running_stats:
            with torch.no_grad():
                self.num_batches_tracked += 1
                m = self.momentum if self.momentum is not None else 1.0 / float(self.num_batches_tracked)
                unbiased = var * (n / (n - 1).clamp_min(1.0))
                self.running_mean.mul_(1 - m).add_(mean.to(self.running_mean.dtype), alpha=m)
                self.running_var.mul_(1 - m).add_(unbiased.to(self.running_var.dtype), alpha=m)
        return y


SynchronizedBatchNorm1d = SynchronizedBatchNorm2d = SynchronizedBatchNorm3d = SynchronizedBatchNorm


def convert_model(module: nn.Module, process_group=None) -> nn.Module:
    """Replace every ``nn.BatchNorm*d`` in ``module`` by :class:`SynchronizedBatchNorm`, keeping parameters and running
    statistics (the reference's ``convert_model``, batchnorm_utils.py:300-340)."""
    if isinstance(module, nn.modules.batchnorm._BatchNorm) and not isinstance(module, SynchronizedBatchNorm):
        new = SynchronizedBatchNorm(module.num_features, module.eps, module.momentum, module.affine, module.track_running_stats,
                                    process_group)
        if module.affine:
            new.weight, new.bias = module.weight, module.bias
        if module.track_running_stats:
            new.running_mean, new.running_var, new.num_batches_tracked = module.running_mean, module.running_var, module.num_batches_tracked
        new.train(module.training)
        return new
    for name, child in list(module.named_children()):
        setattr(module, name, convert_model(child, process_group))
    return module
