"""Cheetah — data-parallel training (``fedml.run_distributed``; a ``pass`` stub in the reference,
P/__init__.py:277-278; closest reference code: DDP in hierarchical cross-silo and examples/centralized/main.py).

One process per GPU.  Parameters and gradients of the replica live in two flat fp32 buckets inside the
peer-visible region (``bind_module_to_flat`` re-points ``p.data`` / ``p.grad`` at views), so after
``loss.backward()`` the whole gradient is already a contiguous buffer every peer can read.  ``step()`` is ONE
launch of ``fb_dp_allreduce_step`` (csrc/cheetah.cu): reduce-scatter by pulling peers' gradient shards over
NVLink, SGD(+momentum, weight decay) on the owned shard, all-gather by pushing the new parameters to every
rank.  No NCCL call and no per-tensor optimizer kernels on the step path.  With ``overlap=True`` the flat gradient is
cut into ``n_buckets`` contiguous buckets (at parameter boundaries); a post-accumulate hook on every parameter counts a
bucket's parameters down and, when the last one has its gradient, launches that bucket's reduce-scatter -> SGD ->
all-gather kernel on a side stream with a capped grid, so the transfer of the late layers' gradients runs under the
backward pass of the early layers; ``step()`` launches whatever is left and joins the side stream.
On CPU / gloo (tests) the same interface falls back to ``all_reduce`` + the same update rule."""
from __future__ import annotations

import ctypes as C
import logging
import os
from typing import List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from .arena import ArenaLayout, bind_module_to_flat


class CheetahEngine:
    def __init__(self, model: nn.Module, device, lr=0.05, momentum=0.0, weight_decay=0.0, nesterov=False, group=None,
                 n_buckets: int = 1, overlap: bool = False, provider="auto"):
        self.device = torch.device(device)
        self.model = model.to(self.device)
        self.lr, self.momentum, self.wd, self.nesterov = lr, momentum, weight_decay, nesterov
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.layout = ArenaLayout(self.model)
        n = self.layout.P_train
        self.n = n
        self.cuda = self.device.type == "cuda"
        self.epoch = 0
        if self.cuda:
            from .. import ops
            from ..ops import native
            from .symmetric import SymmetricRegion
            self.ops, self.native = ops, native
            flag_bytes = ops.round_flag_bytes(64)
            self.region = SymmetricRegion(2 * self.layout.P * 4 + flag_bytes + 4096, self.device, group, provider)
            self.region.carve("param", self.layout.P * 4)
            self.region.carve("grad", self.layout.P * 4)
            self.region.carve("flags", flag_bytes)
            self.flat, self.grad = self.region.tensor("param"), self.region.tensor("grad")
            self.region.tensor("flags", torch.int32).zero_()
            self._peer_param, self._peer_grad = self.region.peer_table("param"), self.region.peer_table("grad")
            self._peer_flags = self.region.peer_table("flags")
        else:
            self.flat = torch.zeros(self.layout.P, dtype=torch.float32)
            self.grad = torch.zeros(self.layout.P, dtype=torch.float32)
        bind_module_to_flat(self.model, self.layout, self.flat, self.grad)
        self.mom = torch.zeros(n, dtype=torch.float32, device=self.device) if momentum else None
        # ---- gradient buckets (contiguous element ranges of the flat arena, cut at parameter boundaries)
        self.overlap = bool(overlap) and self.cuda and self.world > 1
        self.n_buckets = max(1, int(n_buckets)) if self.overlap else 1
        self.max_blocks_overlap = int(os.environ.get("FB_CHEETAH_OVERLAP_BLOCKS", "48"))
        segs = [sg for sg in self.layout.segments if sg.kind == "param"]
        target = (n + self.n_buckets - 1) // self.n_buckets
        self.buckets: List[List[int]] = []          # [offset, length]
        start = 0
        for sg in segs:
            end = sg.offset + (sg.numel + 3) // 4 * 4
            if end - start >= target or sg is segs[-1]:
                self.buckets.append([start, min(end, n) - start])
                start = min(end, n)
        self.buckets = [b for b in self.buckets if b[1] > 0]
        self._bucket_of = {}
        params = dict(self.model.named_parameters())
        self._pending = [0] * len(self.buckets)
        for sg in segs:
            bi = next(i for i, (o, ln) in enumerate(self.buckets) if o <= sg.offset < o + ln)
            self._bucket_of[sg.name] = bi
        self._launched = [False] * len(self.buckets)
        self._side = torch.cuda.Stream(self.device) if self.overlap else None
        if self.overlap:
            for name, p in params.items():
                if p.requires_grad and name in self._bucket_of:
                    p.register_post_accumulate_grad_hook(self._make_hook(self._bucket_of[name]))
            self._arm()
        if self.world > 1:      # replicas start identical
            if self.cuda:
                self.region.barrier()
            dist.broadcast(self.flat, src=0, group=group)
            if self.cuda:
                self.region.barrier()
        self.group = group

    # ---- overlap machinery
    def _arm(self):
        params = dict(self.model.named_parameters())
        self._pending = [0] * len(self.buckets)
        for name, bi in self._bucket_of.items():
            if params[name].requires_grad:
                self._pending[bi] += 1
        self._launched = [False] * len(self.buckets)

    def _make_hook(self, bi):
        def hook(_param):
            self._pending[bi] -= 1
            if self._pending[bi] == 0 and not self._launched[bi]:
                self._side.wait_stream(torch.cuda.current_stream(self.device))     # the bucket's gradients are complete
                with torch.cuda.stream(self._side):
                    self._launch_bucket(bi, self.max_blocks_overlap)
        return hook

    def _launch_bucket(self, bi, max_blocks=0):
        from ..ops.native import FbDpStepParams, check, lib, stream_ptr
        off, ln = self.buckets[bi]
        self.epoch += 1
        p = FbDpStepParams()
        p.peer_grad, p.peer_param, p.peer_flags = self._peer_grad.data_ptr(), self._peer_param.data_ptr(), self._peer_flags.data_ptr()
        p.mom = self.mom.data_ptr() if self.mom is not None else None
        p.n, p.offset, p.rank, p.G, p.epoch = ln, off, self.rank, self.world, self.epoch
        p.lr, p.momentum, p.weight_decay, p.inv_world, p.nesterov = self.lr, self.momentum, self.wd, 1.0 / self.world, int(self.nesterov)
        p.max_blocks = int(max_blocks)
        check(lib().fb_dp_allreduce_step(C.byref(p), C.c_void_p(stream_ptr(self.device))), "dp_allreduce_step")
        self.native.count_launch()
        self._launched[bi] = True

    def zero_grad(self):
        self.grad.zero_()

    def step(self):
        """All-reduce (mean) the gradient bucket and apply SGD on every replica."""
        n = self.n
        if not self.cuda:
            g = self.grad[:n]
            if self.world > 1:
                dist.all_reduce(g, group=self.group)
                g.div_(self.world)
            w = self.flat[:n]
            d = g + self.wd * w if self.wd else g
            if self.mom is not None:
                self.mom.mul_(self.momentum).add_(d)
                d = d + self.momentum * self.mom if self.nesterov else self.mom
            w.sub_(self.lr * d)
            return
        if self.overlap:
            # buckets whose hooks fired are already in flight on the side stream (same order on every rank: autograd visits
            # the parameters of identical replicas in the same order); launch the rest behind them and join
            self._side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self._side):
                for bi in reversed(range(len(self.buckets))):
                    if not self._launched[bi]:
                        self._launch_bucket(bi, 0)
            torch.cuda.current_stream(self.device).wait_stream(self._side)
            self._arm()
            return
        for bi in range(len(self.buckets)):
            self._launch_bucket(bi, 0)

    def state_dict(self):
        return self.layout.unflatten(self.flat.detach().cpu())


class CheetahRunner:
    """``FedMLRunner`` target for ``training_type: distributed``: epochs over this rank's shard of the global
    training set with the fused step."""

    def __init__(self, args, device, dataset, model):
        self.args, self.device = args, torch.device(device)
        world = int(os.environ.get("WORLD_SIZE", "1"))
        if world > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29533")
            dist.init_process_group("nccl" if self.device.type == "cuda" else "gloo", rank=int(os.environ.get("RANK", 0)),
                                    world_size=world)
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.train_global, self.test_global = dataset[2], dataset[3]
        self.engine = CheetahEngine(model, self.device, lr=float(getattr(args, "learning_rate", 0.05)),
                                    momentum=float(getattr(args, "momentum", 0.0) or 0.0),
                                    weight_decay=float(getattr(args, "weight_decay", 0.0) or 0.0))
        self.history = []

    def run(self):
        crit = nn.CrossEntropyLoss()
        m = self.engine.model
        for ep in range(int(getattr(self.args, "epochs", 1))):
            m.train()
            for i, (x, y) in enumerate(self.train_global):
                if i % self.world != self.rank:
                    continue
                self.engine.zero_grad()
                loss = crit(m(x.to(self.device)), y.to(self.device))
                loss.backward()
                self.engine.step()
            m.eval()
            c = n = 0
            with torch.no_grad():
                for x, y in self.test_global:
                    c += int((m(x.to(self.device)).argmax(1).cpu() == y).sum()); n += len(y)
            self.history.append({"epoch": ep, "test_acc": c / max(n, 1), "loss": float(loss.detach())})
        return self.history
