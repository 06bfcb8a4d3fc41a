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


# ------------------------------------------------------------------------------------------------ replicas inside one process
# nn.DataParallel runs the replicas of a module as threads of one process.  The reference synchronises their BatchNorm layers
# with a master / slave message pipe (model/cv/batchnorm_utils.py:20-155, 156-330); the same protocol, written on
# queue.Queue / threading.Condition, with plain differentiable tensor moves in place of ReduceAddCoalesced / Broadcast.
import collections  # noqa: E402
import functools  # noqa: E402
import queue  # noqa: E402
import threading  # noqa: E402

import torch.nn.functional as F  # noqa: E402


class FutureResult:
    """One-slot mailbox: ``put`` once, ``get`` once (blocks until the value is there)."""

    def __init__(self):
        self._result, self._cond = None, threading.Condition()

    def put(self, result):
        with self._cond:
            assert self._result is None, "previous result has not been fetched"
            self._result = result
            self._cond.notify()

    def get(self):
        with self._cond:
            while self._result is None:
                self._cond.wait()
            res, self._result = self._result, None
            return res


_SlavePipeBase = collections.namedtuple("_SlavePipeBase", ["identifier", "queue", "result"])


class SlavePipe(_SlavePipeBase):
    """A replica's end of the pipe: send a message to the master, wait for the reply, acknowledge."""

    def run_slave(self, msg):
        self.queue.put((self.identifier, msg))
        ret = self.result.get()
        self.queue.put(True)
        return ret


class SyncMaster:
    """The master replica's end: collects one message per registered replica, lets ``master_callback`` turn all messages into
    per-replica replies, hands them out and waits for the acknowledgements.  The registry is rebuilt at every replication."""

    def __init__(self, master_callback):
        self._master_callback = master_callback
        self._queue: "queue.Queue" = queue.Queue()
        self._registry: "collections.OrderedDict" = collections.OrderedDict()
        self._activated = False

    def __getstate__(self):
        return {"master_callback": self._master_callback}

    def __setstate__(self, state):
        self.__init__(state["master_callback"])

    def register_slave(self, identifier) -> SlavePipe:
        if self._activated:
            assert self._queue.empty(), "queue is not clean before the next initialisation"
            self._activated = False
            self._registry.clear()
        future = FutureResult()
        self._registry[identifier] = future
        return SlavePipe(identifier, self._queue, future)

    def run_master(self, master_msg):
        self._activated = True
        intermediates = [(0, master_msg)]
        for _ in range(self.nr_slaves):
            intermediates.append(self._queue.get())
        results = self._master_callback(intermediates)
        assert results[0][0] == 0, "the first reply must be the master's"
        for i, res in results:
            if i != 0:
                self._registry[i].put(res)
        for _ in range(self.nr_slaves):
            assert self._queue.get() is True
        return results[0][1]

    @property
    def nr_slaves(self):
        return len(self._registry)


_ChildMessage = collections.namedtuple("_ChildMessage", ["sum", "ssum", "sum_size"])
_MasterMessage = collections.namedtuple("_MasterMessage", ["mean", "inv_std"])


class ReplicaSynchronizedBatchNorm(nn.modules.batchnorm._BatchNorm):
    """BatchNorm whose replicas under :class:`DataParallelWithCallback` normalise with the statistics of the whole batch.
    Outside data-parallel replication (or in eval mode) it is plain BatchNorm.  Accepts 2-D to 5-D inputs."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True):
        super().__init__(num_features, eps=eps, momentum=momentum, affine=affine)
        self._sync_master = SyncMaster(self._data_parallel_master)
        self._is_parallel, self._parallel_id, self._slave_pipe = False, None, None

    def _check_input_dim(self, x):
        if x.dim() < 2:
            raise ValueError(f"expected at least 2D input (got {x.dim()}D input)")

    def forward(self, x):
        if not (self._is_parallel and self.training):
            return F.batch_norm(x, self.running_mean, self.running_var, self.weight, self.bias, self.training, self.momentum, self.eps)
        shape = x.shape
        flat = x.reshape(shape[0], self.num_features, -1)
        msg = _ChildMessage(flat.sum(dim=(0, 2)), (flat * flat).sum(dim=(0, 2)), flat.shape[0] * flat.shape[2])
        mean, inv_std = self._sync_master.run_master(msg) if self._parallel_id == 0 else self._slave_pipe.run_slave(msg)
        out = (flat - mean.view(1, -1, 1)) * inv_std.view(1, -1, 1)
        if self.affine:
            out = out * self.weight.view(1, -1, 1) + self.bias.view(1, -1, 1)
        return out.reshape(shape)

    def __data_parallel_replicate__(self, ctx, copy_id):
        self._is_parallel, self._parallel_id = True, copy_id
        if copy_id == 0:
            ctx.sync_master = self._sync_master
        else:
            self._slave_pipe = ctx.sync_master.register_slave(copy_id)

    def _data_parallel_master(self, intermediates):
        """All replicas' (sum, sum of squares, count) -> (mean, 1/std) for each of them; updates the running statistics."""
        intermediates = sorted(intermediates, key=lambda i: i[0])
        dev = intermediates[0][1].sum.device
        total = sum(i[1].sum_size for i in intermediates)
        s = functools.reduce(torch.add, [i[1].sum.to(dev) for i in intermediates])
        ss = functools.reduce(torch.add, [i[1].ssum.to(dev) for i in intermediates])
        assert total > 1, "BatchNorm computes unbiased standard deviation, which requires size > 1"
        mean = s / total
        sumvar = ss - s * mean
        with torch.no_grad():
            self.running_mean.mul_(1 - self.momentum).add_(self.momentum * mean.detach())
            self.running_var.mul_(1 - self.momentum).add_(self.momentum * sumvar.detach() / (total - 1))
            if self.num_batches_tracked is not None:
                self.num_batches_tracked += 1
        inv_std = (sumvar / total).clamp_min(0.0).add(self.eps).rsqrt()
        return [(i[0], _MasterMessage(mean.to(i[1].sum.device), inv_std.to(i[1].sum.device))) for i in intermediates]


class CallbackContext:
    pass


def execute_replication_callbacks(modules):
    """After ``replicate``: calls ``__data_parallel_replicate__(ctx, copy_id)`` on every sub-module that defines it; the j-th
    sub-module of all replicas shares one context, and the master copy (id 0) is called first."""
    master = modules[0]
    ctxs = [CallbackContext() for _ in master.modules()]
    for i, module in enumerate(modules):
        for j, m in enumerate(module.modules()):
            if hasattr(m, "__data_parallel_replicate__"):
                m.__data_parallel_replicate__(ctxs[j], i)


class DataParallelWithCallback(nn.DataParallel):
    """``nn.DataParallel`` that runs the replication callbacks, so :class:`ReplicaSynchronizedBatchNorm` layers find each other."""

    def replicate(self, module, device_ids):
        modules = super().replicate(module, device_ids)
        execute_replication_callbacks(modules)
        return modules


def patch_replication_callback(data_parallel: nn.DataParallel):
    """Adds the callback step to an existing ``nn.DataParallel`` object."""
    assert isinstance(data_parallel, nn.DataParallel)
    old = data_parallel.replicate

    @functools.wraps(old)
    def new_replicate(module, device_ids):
        modules = old(module, device_ids)
        execute_replication_callbacks(modules)
        return modules

    data_parallel.replicate = new_replicate
