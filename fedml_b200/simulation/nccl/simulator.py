"""The multi-GPU ("NCCL") Parrot simulator — a stub in the reference, the product path here.

Reference state: ``SimulatorNCCL.__init__`` sets ``self.simulator = None`` (simulation/simulator.py:100-108)
and ``simulation/nccl/README.md`` only says "AllReduce-based Federated Learning with Resource
Scheduling".  Semantics are therefore lifted from the SP/MPI FedAvg loops:
per round ``np.random.seed(round)`` client sampling, sample-count weighting over every state entry,
periodic evaluation — identical for any number of GPUs.

Design (one process per GPU, ``torch.distributed`` only for bootstrap):
  * the K sampled clients are assigned to ranks by LPT scheduling on their sample counts
    (core/schedule/scheduler.py semantics);
  * each rank keeps its clients' models as rows of a device-resident fp32 stack ``[C, P]`` (no
    ``state_dict`` copies, no D2H), broadcasts the global model into the rows, and trains all rows
    together with the client-batched executor (models/batched.py) under a CUDA graph — or, for model
    families / custom trainers without a batched implementation, one client at a time on
    arena-bound parameters;
  * inputs come from the packed host dataset through one pinned staging buffer and ONE H2D copy per
    round and rank;
  * the round ends with ONE fused kernel (csrc/round.cu) that weights, clips/noises, reduces across
    GPUs over peer memory, applies the server optimizer and leaves the new global model on every
    rank; robust aggregators (median / trimmed mean / Krum) use csrc/robust.cu on peer rows.
On a CPU-only box the same control flow runs with the PyTorch oracles and a gloo all-reduce, which
is what the multi-process CPU tests exercise.
"""
from __future__ import annotations

import logging
import math
import os
import time
from collections import OrderedDict
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.distributed as dist

from ...core import aggregation as A
from ...core.schedule import schedule_clients
from ...constants import canonical_optimizer_name
from ...models.batched import NativeBackend, TorchBackend, make_batched
from ...ops import native as _native
from ...parallel.arena import ArenaLayout, bind_module_to_flat
from ...utils.metrics import MetricsSink
from ..sp.fedavg_api import client_sampling
from ..trainers import create_model_trainer


def _dist_setup(device: torch.device):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        backend = "nccl" if device.type == "cuda" else "gloo"
        kw = {"device_id": device} if device.type == "cuda" else {}
        dist.init_process_group(backend=backend, rank=int(os.environ.get("RANK", "0")), world_size=world, **kw)
    if dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class PhaseTimer:
    """CUDA-event timers per phase; ``summary`` reduces to the max over ranks (measurement rule of
    BASELINE.json: multi-GPU numbers are device times, max over ranks)."""

    def __init__(self, device):
        self.device, self.cuda = device, device.type == "cuda"
        self.pending: Dict[str, list] = {}
        self.totals: Dict[str, float] = {}

    # NVTX ranges carry the reference's MLOps event names so a timeline reads like its log
    # (core/mlops/mlops_profiler_event.py: "train", "aggregate", "comm_c2s", "server.wait")
    NVTX = os.environ.get("FB_NVTX", "1") != "0"

    def reset(self):
        """Drop everything measured so far (e.g. warm-up rounds that include CUDA-graph capture)."""
        self.collect()
        self.totals = {}

    def start(self, name):
        if self.cuda:
            if self.NVTX:
                torch.cuda.nvtx.range_push(name)
            ev = torch.cuda.Event(enable_timing=True); ev.record()
        else:
            ev = time.perf_counter()
        self.pending.setdefault(name, []).append([ev, None])

    def stop(self, name):
        if self.cuda:
            ev = torch.cuda.Event(enable_timing=True); ev.record()
            if self.NVTX:
                torch.cuda.nvtx.range_pop()
        else:
            ev = time.perf_counter()
        self.pending[name][-1][1] = ev

    def collect(self):
        if self.cuda:
            torch.cuda.synchronize(self.device)
        for name, pairs in self.pending.items():
            for a, b in pairs:
                if b is None:
                    continue
                ms = a.elapsed_time(b) if self.cuda else (b - a) * 1e3
                self.totals[name] = self.totals.get(name, 0.0) + ms
        self.pending = {}
        return dict(self.totals)


class NCCLSimulator:
    def __init__(self, args, device, dataset, model, model_trainer=None, server_aggregator=None):
        self.args = args
        self.device = torch.device(device)
        self.rank, self.world = _dist_setup(self.device)
        self.cuda = self.device.type == "cuda"
        if self.cuda:
            torch.cuda.set_device(self.device)
        ds = list(dataset)
        (self.train_num, self.test_num, self.train_global, self.test_global, self.local_num, self.train_local,
         self.test_local, self.class_num) = ds[:8]
        self.fed = getattr(args, "_federated_dataset", None)
        self.model = model
        self.layout = ArenaLayout(model)
        self.P, self.P_train = self.layout.P, self.layout.P_train
        self.alg = canonical_optimizer_name(getattr(args, "federated_optimizer", "FedAvg"))
        self.robust = str(getattr(args, "robust_agg", "none") or "none").lower()
        dt = str(getattr(args, "defense_type", "") or "").lower()
        if self.alg == "FedAvg_robust" and self.robust == "none" and dt:
            self.robust = {"norm_diff_clipping": "clip", "weak_dp": "weak_dp", "median": "median",
                           "coordinate_median": "median", "trimmed_mean": "trimmed_mean", "krum": "krum",
                           "multi_krum": "krum"}.get(dt, "none")
        self.metrics = MetricsSink(args) if self.rank == 0 else None
        self.timer = PhaseTimer(self.device)
        self.history: List[dict] = []
        self.custom_trainer = model_trainer
        self.K = min(int(args.client_num_per_round), int(args.client_num_in_total))
        self.C_max = int(math.ceil(self.K / self.world))
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self._build_engine()
        self._build_executor()
        self.round_idx = 0

    # ------------------------------------------------------------------ set-up
    def _build_engine(self):
        args = self.args
        flat0 = self.layout.flatten(self.model.state_dict())
        clip = float(getattr(args, "dp_clip", 0.0) or 0.0)
        sigma = float(getattr(args, "dp_sigma", 0.0) or 0.0)
        if self.robust in ("clip", "weak_dp"):
            clip = clip or float(getattr(args, "norm_bound", 5.0))
            if self.robust == "weak_dp":
                sigma = sigma or float(getattr(args, "stddev", 0.025))
        self.dp_clip, self.dp_sigma = clip, sigma
        use_opt = self.alg == "FedOpt"
        if self.cuda:
            from ...parallel.round_engine import DPConfig, RoundEngine, ServerOptConfig
            two_shot = getattr(args, "two_shot", None)
            self.engine = RoundEngine(self.layout, self.device, opt=ServerOptConfig.from_args(args, use_opt),
                                      dp=DPConfig(clip, sigma, int(getattr(args, "random_seed", 0)),
                                                  aggregate_noise=str(getattr(args, "dp_noise_mode", "per_client")) == "aggregate"),
                                      use_multicast=bool(getattr(args, "use_multicast", True)),
                                      two_shot=two_shot, max_clients=max(self.C_max, 1),
                                      provider=str(getattr(args, "symmetric_provider", "auto")))
            self.engine.set_global(flat0)
            self.global_flat = self.engine.global_model
            need_peer_rows = self.robust in ("median", "trimmed_mean", "krum") and self.world > 1
            if need_peer_rows:
                from ...parallel.symmetric import SymmetricRegion
                self.stack_region = SymmetricRegion(self.C_max * self.P * 4 + 4096, self.device,
                                                    provider=self.engine.region.provider)
                self.stack_region.carve("stack", self.C_max * self.P * 4)
                self.stack = self.stack_region.tensor("stack").view(self.C_max, self.P)
            else:
                self.stack_region = None
                self.stack = torch.zeros(self.C_max, self.P, dtype=torch.float32, device=self.device)
        else:
            self.engine = None
            self.global_flat = flat0.clone()
            self.stack = torch.zeros(self.C_max, self.P, dtype=torch.float32)
            self.server_opt = A.ServerOptimizer(
                getattr(args, "server_optimizer", "sgd"), lr=float(getattr(args, "server_lr", 1.0)),
                momentum=float(getattr(args, "server_momentum", 0.0)), beta1=float(getattr(args, "server_beta1", 0.9)),
                beta2=float(getattr(args, "server_beta2", 0.99)), eps=float(getattr(args, "server_eps", 1e-3))) \
                if use_opt else None
        self.grad_stack = torch.zeros_like(self.stack)
        self.W = self.layout.stacked_views(self.stack)
        self.G = self.layout.stacked_views(self.grad_stack)

    def _build_executor(self):
        args = self.args
        self.batched = None
        if int(getattr(args, "client_groups", 0) or os.environ.get("FB_CLIENT_GROUPS", "0") or 0) > 1:
            # concurrent client groups launch BatchNorm kernels on several streams at once; the single-launch form shares
            # one in-library barrier per client index, so keep the two-pass form (read once, before the first BN call)
            os.environ.setdefault("FB_BN_FUSED", "0")
        if self.custom_trainer is None and bool(getattr(args, "batched_clients", True)):
            if self.cuda:
                # "tf32" / "fp32": fp32 activations and weights in HBM, TF32 tensor-core products (the reference's precision)
                dtype = {"bf16": torch.bfloat16, "fp32": torch.float32, "tf32": torch.float32, "fp16": torch.float16}[
                    str(getattr(args, "compute_dtype", "bf16")).lower()]
                # batched_backend "torch": the same client-batched control flow on plain PyTorch fp32 ops (the oracle the
                # multi-GPU tests compare the native path against)
                backend = TorchBackend(torch.float32) if str(getattr(args, "batched_backend", "native")) == "torch" \
                    else NativeBackend(dtype)
            else:
                backend = TorchBackend(torch.float32)
            self.batched = make_batched(self.model, backend)
        if self.batched is None:
            self._ensure_sequential()
        self.lr = float(getattr(args, "learning_rate", getattr(args, "lr", 0.03)))
        self.momentum = float(getattr(args, "momentum", 0.0) or 0.0) if self.alg == "FedNova" or \
            str(getattr(args, "client_optimizer", "sgd")).lower() == "sgd_momentum" else 0.0
        self.wd = float(getattr(args, "wd", 0.0) or 0.0) if self.alg == "FedNova" else 0.0
        self.mu = float(getattr(args, "fedprox_mu", 0.0) or getattr(args, "mu", 0.0) or 0.0) \
            if self.alg in ("FedProx", "FedNova") else 0.0
        if self.alg == "FedProx" and self.mu == 0.0:
            self.mu = 0.1
        self.mom_stack = torch.zeros_like(self.stack) if self.momentum else None
        self.graph = None
        self.use_graph = self.cuda and bool(getattr(args, "use_cuda_graphs", True)) and self.batched is not None
        self._static = None
        self.global_momentum = None   # FedNova gmf buffer (server side)

    # ------------------------------------------------------------------ data staging
    def _client_arrays(self, cid):
        if self.fed is not None:
            return self.fed.client_slice(cid)
        batches = self.train_local[cid]
        return torch.cat([b[0] for b in batches]), torch.cat([b[1] for b in batches])

    def _stage_round_inputs(self, my_clients: List[int], round_idx: int):
        """Inputs of one round for this rank: X [C, S, B, ...], Y [C, S, B] (-1 = padding), steps [C].

        Host side: ONE vectorised gather (``index_select`` with a [C*S*B] index built from the packed
        dataset's client offsets) straight into a persistent pinned staging buffer, then ONE async H2D
        copy on a side stream.  Two staging slots alternate so the next round can be staged while the
        current one trains (see ``_prefetch``)."""
        args = self.args
        bs = int(args.batch_size)
        C = self.C_max
        if self.fed is None:
            return self._stage_round_inputs_generic(my_clients, round_idx)
        dl = getattr(args, "drop_last", None)
        drop_last = (self.fed.transform is not None if dl is None else bool(dl)) and bs > 0
        offs = self.fed.client_offsets
        plans = []
        for cid in my_clients:
            n = int(offs[cid + 1] - offs[cid])
            b = n if bs <= 0 else bs
            steps = n // b if drop_last else (n + b - 1) // b
            plans.append((cid, n, b, steps))
        B = max([p[2] for p in plans], default=max(bs, 1))
        S = max([p[3] for p in plans], default=0)
        if not plans or S == 0:
            return None
        idx = torch.zeros(C, S * B, dtype=torch.int64)
        valid = torch.zeros(C, S, B, dtype=torch.bool)
        steps_t = torch.zeros(C, dtype=torch.int64)
        shuffle = drop_last and bool(getattr(args, "shuffle_clients", True))
        for pos, (cid, n, b, steps) in enumerate(plans):
            if shuffle:
                g = torch.Generator().manual_seed(int(getattr(args, "random_seed", 0)) * 1000003 + round_idx * 1009 + cid)
                local = torch.randperm(n, generator=g)
            else:
                local = torch.arange(n)
            take = min(n, steps * b)
            # lay the client's samples out step-major with B slots per step (ragged last batch stays padded)
            for s_ in range(steps):
                lo, hi = s_ * b, min((s_ + 1) * b, take)
                idx[pos, s_ * B:s_ * B + (hi - lo)] = local[lo:hi] + int(offs[cid])
                valid[pos, s_, :hi - lo] = True
            steps_t[pos] = steps
        slot = self._stage_slot = (getattr(self, "_stage_slot", 1) + 1) % 2
        nrow = C * S * B
        xs, ys = self.fed.x, self.fed.y
        key = (nrow, tuple(xs.shape[1:]), xs.dtype, tuple(ys.shape[1:]))
        if getattr(self, "_pinned_key", None) != key:
            pin = self.cuda
            self._pinned = [(torch.empty((nrow,) + tuple(xs.shape[1:]), dtype=xs.dtype, pin_memory=pin),
                             torch.empty((nrow,) + tuple(ys.shape[1:]), dtype=ys.dtype, pin_memory=pin)) for _ in range(2)]
            self._pinned_key = key
            self._pinned_events = [None, None]
        Xp, Yp = self._pinned[slot]
        if self.cuda and self._pinned_events[slot] is not None:
            self._pinned_events[slot].synchronize()        # the previous H2D out of this slot has finished
        flat = idx.view(-1)
        torch.index_select(xs, 0, flat, out=Xp)
        torch.index_select(ys, 0, flat, out=Yp)
        Yp.view(C, S, B, *ys.shape[1:])[~valid] = -1
        inv_valid = 1.0 / valid.sum(dim=2).clamp_min(1).to(torch.float32)
        # a step that is active for a client but not full: its padding slots would enter BatchNorm batch statistics
        nvalid = valid.sum(dim=2)
        self._ragged_round = bool(((nvalid > 0) & (nvalid < B)).any())
        if self.cuda:
            if getattr(self, "_copy_stream", None) is None:
                self._copy_stream = torch.cuda.Stream(self.device)
            with torch.cuda.stream(self._copy_stream):
                Xd = Xp.to(self.device, non_blocking=True).view(C, S, B, *xs.shape[1:])
                Yd = Yp.to(self.device, non_blocking=True).view(C, S, B, *ys.shape[1:])
                inv_d = inv_valid.to(self.device, non_blocking=True)
                steps_d = steps_t.to(self.device, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self._copy_stream)
            self._pinned_events[slot] = ev
            self.h2d_bytes += Xp.numel() * Xp.element_size() + Yp.numel() * Yp.element_size()
        else:
            Xd, Yd = Xp.view(C, S, B, *xs.shape[1:]).clone(), Yp.view(C, S, B, *ys.shape[1:]).clone()
            inv_d, steps_d, ev = inv_valid, steps_t, None
        return Xd, Yd, steps_d, inv_d, S, ev

    def _stage_round_inputs_generic(self, my_clients, round_idx):
        """Fallback for datasets handed over only as the 8-tuple (lists of batches / DataLoaders)."""
        C = self.C_max
        per = []
        for cid in my_clients:
            batches = list(self.train_local[cid])
            per.append(batches)
        S = max((len(b) for b in per), default=0)
        if S == 0:
            return None
        B = max(b[0].shape[0] for bl in per for b in bl)
        x0, y0 = per[0][0]
        X = torch.zeros((C, S, B) + tuple(x0.shape[1:]), dtype=x0.dtype)
        Y = torch.full((C, S, B) + tuple(y0.shape[1:]), -1, dtype=torch.int64)
        steps_t = torch.zeros(C, dtype=torch.int64)
        valid = torch.ones(C, S)
        for pos, bl in enumerate(per):
            steps_t[pos] = len(bl)
            for s_, (xb, yb) in enumerate(bl):
                X[pos, s_, :xb.shape[0]] = xb
                Y[pos, s_, :yb.shape[0]] = yb
                valid[pos, s_] = xb.shape[0]
        dev = self.device
        if self.cuda:
            self.h2d_bytes += X.numel() * X.element_size() + Y.numel() * Y.element_size()
        return X.to(dev), Y.to(dev), steps_t.to(dev), (1.0 / valid).to(dev), S, None

    def _round_plan(self, round_idx: int):
        N = int(self.args.client_num_in_total)
        sampled = client_sampling(round_idx, N, self.K)
        sizes = [int(self.local_num[c]) for c in sampled]
        assign = schedule_clients(sizes, self.world)
        my_pos = assign[self.rank]
        return sampled, sizes, assign, my_pos, [sampled[p] for p in my_pos]

    def _prefetch(self, round_idx: int):
        """Stage the inputs of a future round while the GPU is still busy with the current one."""
        if self.batched is None or round_idx >= int(self.args.comm_round):
            return
        plan = self._round_plan(round_idx)
        self._prefetched = (round_idx, plan, self._stage_round_inputs(plan[4], round_idx) if plan[4] else None)

    # ------------------------------------------------------------------ local training
    def _sparse_sgd_range(self):
        """[a, b) of the arena whose update the batched executor applies itself in sparse form (word embeddings under
        plain SGD: only the rows of tokens that occur are touched), or None."""
        if not hasattr(self, "_sparse_range"):
            rng = None
            name = getattr(self.batched, "sparse_sgd_segment", None)
            plain = not (self.momentum or self.wd or self.mu)
            if name and plain and bool(getattr(self.args, "sparse_embedding_sgd", True)):
                seg = self.layout.by_name[name]
                if seg.offset % 4 == 0 and (seg.offset + seg.numel) % 4 == 0:
                    rng = (seg.offset, seg.offset + seg.numel)
            self._sparse_range = rng
        return self._sparse_range

    def _sgd(self, C, lr_vec, first_step, rows=None):
        """Fused multi-client optimizer step on arena rows [r0, r1) (default: the first C rows)."""
        n = self.P_train
        r0, r1 = rows if rows is not None else (0, C)
        anchor = self.global_flat if self.mu else None
        skip = self._sparse_sgd_range() if self.batched is not None else None
        if skip is not None:
            # dense optimizer step on the two sides of the sparsely-updated segment
            for a, b in ((0, skip[0]), (skip[1], n)):
                if b <= a:
                    continue
                if self.cuda:
                    from ... import ops
                    ops.sgd_step_stacked(self.stack[r0:r1, a:b], self.grad_stack[r0:r1, a:b], b - a, lr_vec)
                else:
                    self.stack[r0:r1, a:b].sub_(lr_vec[:r1 - r0, None] * self.grad_stack[r0:r1, a:b])
            return
        if self.cuda:
            from ... import ops
            ops.sgd_step_stacked(self.stack[r0:r1], self.grad_stack[r0:r1], n, lr_vec, weight_decay=self.wd,
                                 momentum=self.momentum, mom_buf=None if self.mom_stack is None else self.mom_stack[r0:r1],
                                 first_step=first_step, mu=self.mu, anchor=anchor)
        else:
            C = r1 - r0
            w, g = self.stack[r0:r1, :n], self.grad_stack[r0:r1, :n]
            d = g + self.wd * w if self.wd else g.clone()
            if self.momentum:
                buf = self.mom_stack[r0:r1, :n]
                if first_step:
                    buf.copy_(d)
                else:
                    buf.mul_(self.momentum).add_(d)
                d = buf.clone()
            if self.mu:
                d = d + self.mu * (w - anchor[:n])
            w.sub_(lr_vec[:C, None] * d)

    def _ensure_sequential(self):
        """Per-client eager trainer on arena-bound parameters (families without a batched executor, custom trainers, and
        rounds the batched executor must not take)."""
        if getattr(self, "seq_trainer", None) is None:
            import copy
            # the fallback of a batched executor gets its own module; otherwise bind the caller's module (a custom trainer
            # trains that very object)
            self.seq_model = (copy.deepcopy(self.model) if getattr(self, "batched", None) is not None else self.model).to(self.device)
            self.seq_flat = torch.zeros(self.P, dtype=torch.float32, device=self.device)
            bind_module_to_flat(self.seq_model, self.layout, self.seq_flat)
            self.seq_trainer = self.custom_trainer or create_model_trainer(self.seq_model, self.args)

    def _client_groups(self):
        """Optional (``client_groups`` > 1): split the clients of this GPU into groups that advance on separate
        streams; inside the captured CUDA graph these are parallel branches.  Each group has its own executor
        (staging buffers, side stream, scratch).  Off by default, see the measurement note below."""
        if not hasattr(self, "_groups"):
            C = self.C_max
            want = int(getattr(self.args, "client_groups", 0) or os.environ.get("FB_CLIENT_GROUPS", "0") or 0)
            if want <= 0:
                # measured on B200 (ResNet-56, 13 / 25 / 50 clients): 2-4 groups are 1-8 % SLOWER than one group, because
                # every kernel already sizes its grid to the whole GPU; kept as an option for models with tiny layers
                want = 1
            want = max(1, min(want, C))
            groups = []
            if self.cuda and want > 1 and not self.momentum:
                from ...models.batched import NativeBackend, make_batched
                bounds = [round(i * C / want) for i in range(want + 1)]
                for g in range(want):
                    a, b = bounds[g], bounds[g + 1]
                    if b <= a:
                        continue
                    ex = make_batched(self.model, NativeBackend(self.batched.be.dtype))
                    groups.append({"rows": (a, b), "exec": ex, "stream": torch.cuda.Stream(self.device),
                                   "W": self.layout.stacked_views(self.stack[a:b]),
                                   "G": self.layout.stacked_views(self.grad_stack[a:b])})
            self._groups = groups
        return self._groups

    def _step_body(self, st):
        sparse = self._sparse_sgd_range() is not None
        groups = self._client_groups()
        if groups:
            cur = torch.cuda.current_stream()
            for gr in groups:
                a, b = gr["rows"]
                gr["stream"].wait_stream(cur)
                with torch.cuda.stream(gr["stream"]):
                    if sparse:
                        gr["exec"].set_sparse_sgd(st["lr"][a:b])
                    loss = gr["exec"].train_step(st["x"][a:b], st["y"][a:b], gr["W"], gr["G"], st["active"][a:b],
                                                 st["scale"][a:b])
                    self._sgd(b - a, st["lr"][a:b], False, rows=(a, b))
                    st["loss"][a:b].copy_(loss)
            for gr in groups:
                cur.wait_stream(gr["stream"])
            return
        if sparse:
            self.batched.set_sparse_sgd(st["lr"])
        fused = self._fused_sgd()
        if fused:
            self.batched.set_fused_sgd(st["lr"])
        x_in = st["x"]
        if st.get("aug_nonce") is not None and x_in.dtype == torch.uint8 and x_in.dim() == 5:
            # train-time crop / flip / Cutout inside the step (and inside the captured graph): the nonce lives on the
            # device and is bumped by the host copy before every step, so each replay draws new transforms
            from ...data.augment import augment_cifar
            aug = self.fed.train_augment
            x_in = augment_cifar(x_in, aug.seed, st["aug_nonce"], aug.pad, aug.cutout, aug.mean_u8)
        loss = self.batched.train_step(x_in, st["y"], self.W, self.G, st["active"], st["scale"])
        if not (fused and getattr(self.batched, "fused_applied", False)):
            self._sgd(self.C_max, st["lr"], False)
        st["loss"].copy_(loss)

    def _fused_sgd(self):
        """Plain SGD on the native back-end: the executor applies the optimizer step itself, fused with re-staging the
        next step's bf16 operands (``fb_sgd_restage``)."""
        if not hasattr(self, "_fused_ok"):
            plain = not (self.momentum or self.wd or self.mu)
            self._fused_ok = bool(self.cuda and plain and getattr(self.batched, "supports_fused_sgd", False)
                                  and isinstance(getattr(self.batched, "be", None), NativeBackend)
                                  and bool(getattr(self.args, "fused_sgd_restage", True))
                                  and os.environ.get("FB_FUSED_SGD", "1") != "0")
        return self._fused_ok

    def _train_batched(self, staged, n_local):
        Xd, Yd, steps_dev, inv_valid, S, ev = staged
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)       # inputs were copied on the side stream
        C = self.C_max
        epochs = int(getattr(self.args, "epochs", 1))
        dev = self.device
        if self._static is None or self._static["x"].shape != Xd[:, 0].shape:
            self._static = {"x": torch.empty_like(Xd[:, 0]), "y": torch.empty_like(Yd[:, 0]),
                            "active": torch.zeros(C, dtype=torch.float32, device=dev),
                            "lr": torch.zeros(C, dtype=torch.float32, device=dev),
                            "scale": torch.zeros(C, dtype=torch.float32, device=dev),
                            "loss": torch.zeros(C, dtype=torch.float32, device=dev)}
            if self.fed is not None and getattr(self.fed, "train_augment", None) is not None and self.cuda:
                self._static["aug_nonce"] = torch.zeros(1, dtype=torch.int64, device=dev)
                self._aug_count = 0
            self.graph = None
        st = self._static
        first = True
        if self._fused_sgd():
            self.batched.restage(self.W)                     # the arena holds the fresh global model: stage it once
        for _ in range(epochs):
            for s in range(S):
                act = (steps_dev > s).to(torch.float32)
                st["x"].copy_(Xd[:, s]); st["y"].copy_(Yd[:, s])
                st["active"].copy_(act); st["lr"].copy_(act * self.lr); st["scale"].copy_(inv_valid[:, s])
                if st.get("aug_nonce") is not None:
                    self._aug_count += 1
                    # rank-independent nonce; the sample index inside a step is the slot, which is fine for augmentation
                    st["aug_nonce"].fill_(self._aug_count * 64 + self.rank)
                if self.momentum and first:
                    loss = self.batched.train_step(st["x"], st["y"], self.W, self.G, st["active"], st["scale"])
                    self._sgd(C, st["lr"], True)
                    st["loss"].copy_(loss)
                elif self.use_graph:
                    if self.graph is None:
                        self._capture(st)
                    self.graph.replay()
                    _native.count_launch(self._graph_launches)   # kernels of ours inside the replayed graph
                else:
                    self._step_body(st)
                first = False
        if self._fused_sgd():
            self.batched.invalidate_stage()                  # aggregation rewrites the arena behind the executor's back
        return st["loss"]

    def _capture(self, st):
        # warm-up on a side stream (allocator + lazy kernel loading), then capture one step
        backup = self.stack.clone()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                self._step_body(st)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.stack.copy_(backup)          # warm-up steps must not advance the models
        if self._fused_sgd():
            self.batched.restage(self.W)
        self.graph = torch.cuda.CUDAGraph()
        before = _native.LAUNCHES[0]
        with torch.cuda.graph(self.graph):
            self._step_body(st)
        self._graph_launches = _native.LAUNCHES[0] - before
        _native.LAUNCHES[0] = before
        self.stack.copy_(backup)          # capture does not execute, but keep the invariant explicit
        if self._fused_sgd():
            self.batched.restage(self.W)

    def _train_sequential(self, my_clients, round_idx):
        """Per-client eager training on arena-bound parameters (custom trainers / unbatched families)."""
        int_segs = [sg for sg in self.layout.segments if sg.kind == "int_buffer"]
        bufs = dict(self.seq_model.named_buffers())
        for pos, cid in enumerate(my_clients):
            self.seq_flat.copy_(self.global_flat)
            for sg in int_segs:       # integer buffers (num_batches_tracked) are not views of the arena: sync by hand
                bufs[sg.name].copy_(self.global_flat[sg.offset:sg.offset + sg.numel].view(sg.shape).round().to(sg.dtype))
            self.seq_trainer.set_id(cid)
            self.seq_trainer.train(self.train_local[cid], self.device, self.args)
            for sg in int_segs:
                self.seq_flat[sg.offset:sg.offset + sg.numel] = bufs[sg.name].reshape(-1).to(torch.float32)
            self.stack[pos].copy_(self.seq_flat)

    # ------------------------------------------------------------------ one round
    def train_one_round(self, round_idx: int):
        args = self.args
        pre = getattr(self, "_prefetched", None)
        t = self.timer
        t.start("h2d")
        if pre is not None and pre[0] == round_idx:
            (sampled, sizes, assign, my_pos, my_clients), staged = pre[1], pre[2]
        else:
            sampled, sizes, assign, my_pos, my_clients = self._round_plan(round_idx)
            staged = self._stage_round_inputs(my_clients, round_idx) if self.batched is not None and my_clients else None
        self._prefetched = None
        t.stop("h2d")
        dev = self.device
        # every row starts from the global model (dummy rows too: they carry weight 0)
        t.start("train")
        if self.cuda:
            from ... import ops
            ops.broadcast_rows(self.global_flat, self.stack)
        else:
            self.stack.copy_(self.global_flat[None, :].expand_as(self.stack))
        use_batched = self.batched is not None
        if use_batched and staged is not None and getattr(self, "_ragged_round", False) \
                and getattr(self.batched, "uses_batch_statistics", False):
            # BatchNorm normalises over the rows of a batch: the padded slots of a ragged last batch (n % batch_size != 0
            # without drop_last) would change mean / variance and the gradients.  Such rounds train client by client,
            # exactly as the reference does (my_model_trainer_classification.py:21-70).
            if not getattr(self, "_warned_ragged", False):
                logging.warning("ragged batches with a BatchNorm model: this round trains sequentially (set drop_last / a "
                                "batch size dividing the client sizes to use the client-batched kernels)")
                self._warned_ragged = True
            use_batched = False
            self._ensure_sequential()
        if use_batched:
            if staged is not None:
                self._train_batched(staged, sizes)
        else:
            self._train_sequential(my_clients, round_idx)
        t.stop("train")
        # ---- aggregation
        t.start("aggregate")
        weights = torch.zeros(self.C_max, dtype=torch.float32)
        for i, p in enumerate(my_pos):
            weights[i] = sizes[p]
        total = float(sum(sizes))
        self._aggregate(weights.to(dev, non_blocking=True), total, my_pos, sizes, round_idx)
        t.stop("aggregate")
        self.round_idx = round_idx + 1
        # the GPU is still working through this round: use the host time to stage the next one
        if bool(getattr(args, "prefetch_inputs", True)):
            self._prefetch(round_idx + 1)

    def _fednova_coefs(self, sizes, my_pos):
        """Host-side FedNova scalars (fednova.py:114-167, client.py:50-65): lockstep local steps make
        a_i a closed form of (steps, momentum, lr*mu)."""
        args = self.args
        bs = int(args.batch_size)
        total = float(sum(sizes))
        coefs, taus = [], []
        for n in sizes:
            steps = (n // bs if (self.fed is not None and self.fed.transform is not None) else math.ceil(n / max(bs, 1))) \
                * int(getattr(args, "epochs", 1)) if bs > 0 else int(getattr(args, "epochs", 1))
            a, cnt = 0.0, 0.0
            etamu = self.lr * self.mu
            for _ in range(max(steps, 0)):
                if self.momentum:
                    cnt = cnt * self.momentum + 1.0
                    a += cnt
                if etamu:
                    a = a * (1.0 - etamu) + 1.0
                if not self.momentum and not etamu:
                    a += 1.0
            ratio = n / total
            tau = steps * ratio if self.mu != 0 else a * ratio
            coefs.append(ratio / a if a else 0.0)
            taus.append(tau)
        tau_eff = sum(taus)
        return [c * tau_eff for c in coefs]

    def _aggregate(self, weights, total, my_pos, sizes, round_idx):
        args = self.args
        base = sum(len(a) for a in schedule_clients(sizes, self.world)[:self.rank])
        if self.robust in ("median", "trimmed_mean", "krum"):
            return self._aggregate_robust(weights, total, my_pos, sizes, round_idx)
        if bool(getattr(args, "secure_aggregation", False)):
            return self._aggregate_secure(weights, total, my_pos, sizes, round_idx, base)
        if self.alg == "FedNova":
            coefs = self._fednova_coefs(sizes, my_pos)
            cw = torch.zeros(self.C_max, dtype=torch.float32)
            for i, p in enumerate(my_pos):
                cw[i] = coefs[p]
            self_coef = 1.0 - float(sum(coefs))
            gmf = float(getattr(args, "gmf", 0.0) or 0.0)
            if self.cuda:
                eng = self.engine
                if gmf and eng.opt.name != "sgd":
                    from ...parallel.round_engine import ServerOptConfig
                    eng.opt = ServerOptConfig("sgd", lr=self.lr, momentum=gmf)
                    eng.m = torch.zeros(self.P_train, dtype=torch.float32, device=self.device)
                eng.aggregate(self.stack, cw.to(self.device), 1.0, base, round_idx, weights_buf=weights,
                              inv_total_buf=1.0 / total, self_coef=self_coef,
                              grad_scale=(1.0 / self.lr) if gmf else 1.0, use_opt=bool(gmf))
            else:
                part = (cw[:, None] * self.stack).sum(0)
                part_b = (weights[:, None] * self.stack).sum(0) / total
                if self.world > 1:
                    dist.all_reduce(part); dist.all_reduce(part_b)
                Pt = self.P_train
                avg = part[:Pt] + self_coef * self.global_flat[:Pt]
                if gmf:
                    g = (self.global_flat[:Pt] - avg) / self.lr
                    self.global_momentum = g if self.global_momentum is None else self.global_momentum * gmf + g
                    avg = self.global_flat[:Pt] - self.lr * self.global_momentum
                self.global_flat[:Pt] = avg
                self.global_flat[Pt:] = part_b[Pt:]
            return
        if self.cuda:
            self.engine.aggregate(self.stack, weights, 1.0 / total, base, round_idx,
                                  weight_sq_sum=float(sum(float(n) * float(n) for n in sizes)))
            return
        # CPU oracle path (gloo)
        stack = self.stack
        if self.dp_clip > 0 or self.dp_sigma > 0:
            Pt = self.P_train
            diff = stack[:, :Pt] - self.global_flat[:Pt]
            if self.dp_clip > 0:
                diff = diff / torch.clamp(diff.norm(dim=1) / self.dp_clip, min=1.0)[:, None]
            contrib = stack.clone()
            contrib[:, :Pt] = self.global_flat[:Pt] + diff
            if self.dp_sigma > 0:
                g = torch.Generator().manual_seed(int(getattr(args, "random_seed", 0)) * 7919 + round_idx * 131 + self.rank)
                contrib[:, :Pt] += self.dp_sigma * torch.randn(contrib[:, :Pt].shape, generator=g)
            stack = contrib
        part = (weights[:, None] * stack).sum(0) / total
        if self.world > 1:
            dist.all_reduce(part)
        if self.server_opt is not None:
            Pt = self.P_train
            new = self.server_opt.step({"p": self.global_flat[:Pt]}, {"p": part[:Pt]}, ["p"])["p"]
            self.global_flat[:Pt] = new
            self.global_flat[Pt:] = part[Pt:]
        else:
            self.global_flat.copy_(part)

    def _aggregate_secure(self, weights, total, my_pos, sizes, round_idx, base):
        """``secure_aggregation: true``: every simulated client uploads its weighted model as a quantised field vector
        under pairwise masks (core/security/secagg.py; masks expanded by ``fb_ff_mask`` from Philox, never materialised),
        the "server" only ever adds masked vectors (``fb_ff_add``), across GPUs as an integer all-reduce; the pairwise
        masks cancel in the sum, which de-quantises to the weighted mean.  Key agreement runs once per simulator (public
        keys only are shared); the round index enters the mask PRG."""
        from ...core.security.mpc_function import DEFAULT_PRIME
        from ...core.security.secagg import SecAggClient
        K = self.K
        p = DEFAULT_PRIME
        scale = float(getattr(self.args, "secagg_scale", 2 ** 16))
        if getattr(self, "_sec_clients", None) is None:
            rng = np.random.default_rng(int(getattr(self.args, "random_seed", 0)) + 99173)   # same slot keys on every rank
            self._sec_clients = [SecAggClient(i, p, rng=rng) for i in range(K)]
            pks = {c.cid: c.pk for c in self._sec_clients}
            for c in self._sec_clients:
                c.agree(pks)
        parts = list(range(K))
        Pq = self.P
        acc = None
        for i, pos in enumerate(my_pos):
            cl = self._sec_clients[pos]                       # slot `pos` of this round's sample
            row = self.stack[i]
            masked = cl.mask(row, float(sizes[pos]) / total, scale, parts, round_idx=round_idx)
            if acc is None:
                acc = masked.clone()
            elif self.cuda:
                from ... import ops
                ops.ff_add_(acc, masked, p)
            else:
                acc = torch.remainder(acc + masked, p)
        if acc is None:
            acc = torch.zeros(Pq, dtype=torch.int32 if self.cuda else torch.int64, device=self.device)
        if self.world > 1:
            wide = acc.to(torch.int64)
            dist.all_reduce(wide)                              # sum of G residues < G * p: no overflow in int64
            acc = torch.remainder(wide, p).to(acc.dtype)
        if self.cuda:
            from ... import ops
            mean = ops.ff_dequantize(acc, scale, p)
        else:
            signed = torch.where(acc > p // 2, acc - p, acc)
            mean = (signed.double() / scale).to(torch.float32)
        self.global_flat.copy_(mean.view(-1)[:self.P])

    def _aggregate_robust(self, weights, total, my_pos, sizes, round_idx):
        """Median / trimmed mean / Krum over all K client rows (they live on G GPUs)."""
        args = self.args
        K = len(sizes)
        trim = int(math.floor(float(getattr(args, "trim_ratio", 0.1)) * K))
        f = int(getattr(args, "krum_f", 0) or round(float(getattr(args, "byzantine_fraction", 0.0)) * K))
        m = int(getattr(args, "krum_m", 1) or 1)
        self._inject_byzantine(my_pos, K, round_idx)
        if not self.cuda:
            rows = self.stack[:len(my_pos)]
            if self.world > 1:
                gathered = [None] * self.world
                dist.all_gather_object(gathered, (my_pos, rows))
                allrows = torch.zeros(K, self.P)
                for pos, r in gathered:
                    for i, p in enumerate(pos):
                        allrows[p] = r[i]
            else:
                allrows = torch.zeros(K, self.P)
                for i, p in enumerate(my_pos):
                    allrows[p] = rows[i]
            if self.robust == "median":
                self.global_flat.copy_(torch.median(allrows, dim=0).values)
            elif self.robust == "trimmed_mean":
                srt = torch.sort(allrows, dim=0).values
                self.global_flat.copy_(srt[trim:K - trim].mean(0))
            else:
                scores = A.krum_scores(allrows[:, :self.P_train], f)
                chosen = A.krum_select(scores, m)
                w = torch.tensor([float(sizes[i]) for i in chosen])
                self.global_flat.copy_((w[:, None] * allrows[chosen]).sum(0) / w.sum())
            return
        from ...parallel import robust_engine
        robust_engine.aggregate(self, self.robust, K, trim, f, m, my_pos, sizes, round_idx)

    def _inject_byzantine(self, my_pos, K, round_idx):
        frac = float(getattr(self.args, "byzantine_fraction", 0.0) or 0.0)
        n_bad = int(round(frac * K))
        if not n_bad:
            return
        kind = str(getattr(self.args, "byzantine_attack", "sign_flip"))
        Pt = self.P_train
        for i, p in enumerate(my_pos):
            if p < n_bad:   # the first n_bad sampled positions are malicious (deterministic, G-invariant)
                row, g = self.stack[i, :Pt], self.global_flat[:Pt]
                if kind == "sign_flip":
                    row.copy_(g - (row - g))
                elif kind == "scaled":
                    row.copy_(g + 10.0 * (row - g))
                elif kind == "gaussian":
                    gen = torch.Generator(device=row.device).manual_seed(round_idx * 7919 + p)
                    row.copy_(g + torch.randn(row.shape, generator=gen, device=row.device))
                elif kind == "zero":
                    row.zero_()

    # ------------------------------------------------------------------ evaluation
    @torch.no_grad()
    def evaluate(self, round_idx: int):
        """Global-model accuracy/loss over every client's train and test split (SP semantics,
        fedavg_api.py:238-326).  All clients share the global model at this point, so the per-client
        loop of the reference collapses into batched inference over the packed arrays."""
        stats = {"round": round_idx}
        if self.fed is None:
            return stats
        if self.cuda and self.batched is not None and bool(getattr(self.args, "native_eval", True)) \
                and isinstance(getattr(self.batched, "be", None), NativeBackend):
            try:
                return self._evaluate_native(dict(stats), round_idx)
            except (NotImplementedError, TypeError) as exc:      # executor without an eval forward for this input kind
                logging.info("native evaluation not available for this model (%s): eager evaluation", exc)
        sd = self.layout.unflatten(self.global_flat)
        model = self.model.to(self.device)
        model.load_state_dict(sd)
        model.eval()
        cap = int(getattr(self.args, "eval_sample_cap", 0) or 0)
        for split, (x, y) in (("train", (self.fed.x, self.fed.y)), ("test", (self.fed.x_test, self.fed.y_test))):
            n = x.shape[0] if not cap else min(cap, x.shape[0])
            lo = (n * self.rank) // self.world
            hi = (n * (self.rank + 1)) // self.world
            correct = torch.zeros((), device=self.device); loss = torch.zeros((), device=self.device)
            bs = 2048
            for i in range(lo, hi, bs):
                xb, yb = x[i:min(i + bs, hi)], y[i:min(i + bs, hi)]
                xb = self.fed._prep(xb).to(self.device, non_blocking=True)
                yb = yb.to(self.device, non_blocking=True)
                out = model(xb)
                loss += torch.nn.functional.cross_entropy(out, yb, reduction="sum")
                correct += (out.argmax(1) == yb).sum()
            vec = torch.stack([correct, loss, torch.tensor(float(hi - lo), device=self.device)])
            if self.world > 1:
                dist.all_reduce(vec)
            stats[f"{split}_acc"] = float(vec[0] / vec[2].clamp_min(1))
            stats[f"{split}_loss"] = float(vec[1] / vec[2].clamp_min(1))
        self.history.append(stats)
        if self.metrics is not None:
            self.metrics.log({"Train/Acc": stats["train_acc"], "Train/Loss": stats["train_loss"],
                              "Test/Acc": stats["test_acc"], "Test/Loss": stats["test_loss"], "round": round_idx})
        return stats

    def _evaluate_native(self, stats, round_idx):
        """Evaluation on the native kernels: the client-batched executor's eval-mode forward over the packed arrays (one
        "client" = the global model, read in place from the round engine's buffer) and the fused counter kernel
        ``fb_eval_counts`` (K15: arg-max, correct / total, per-class TP / FP / FN and the summed loss in one pass over the
        logits, no host sync per batch).  Replaces my_model_trainer_classification.py:95-154 (per-client eager test loop)."""
        from ... import ops
        from ...models.batched import make_batched
        ex = getattr(self, "_eval_exec", None)
        if ex is None:
            ex = self._eval_exec = make_batched(self.model, NativeBackend(self.batched.be.dtype))
            if ex is None:
                raise NotImplementedError
        g = self.global_flat.view(1, -1)
        W1 = self.layout.stacked_views(g)
        if hasattr(ex, "invalidate_stage"):
            ex.invalidate_stage()
        ncls = int(self.class_num)
        pre_sig = bool(getattr(ex, "pre_sigmoid", False))
        cap = int(getattr(self.args, "eval_sample_cap", 0) or 0)
        bs = int(getattr(self.args, "eval_batch_size", 1024))
        for split, (x, y) in (("train", (self.fed.x, self.fed.y)), ("test", (self.fed.x_test, self.fed.y_test))):
            n = x.shape[0] if not cap else min(cap, x.shape[0])
            lo, hi = (n * self.rank) // self.world, (n * (self.rank + 1)) // self.world
            counts = torch.zeros(2 + 3 * ncls, dtype=torch.int32, device=self.device)
            loss_sum = torch.zeros(1, dtype=torch.float32, device=self.device)
            for i in range(lo, hi, bs):
                xb = x[i:min(i + bs, hi)].to(self.device, non_blocking=True)
                yb = y[i:min(i + bs, hi)].to(self.device, non_blocking=True).to(torch.int64)
                if ex.input_kind != "image_u8_nhwc":
                    xb = self.fed._prep(xb) if xb.dtype == torch.uint8 else xb
                logits = ex.logits(xb.unsqueeze(0), W1)
                ops.eval_counts(logits.reshape(-1, ncls).float().contiguous(), yb.reshape(-1).contiguous(), counts, loss_sum,
                                pre_sigmoid=pre_sig)
            vec = torch.cat([counts.to(torch.float64), loss_sum.to(torch.float64)])
            if self.world > 1:
                dist.all_reduce(vec)
            vec = vec.cpu()
            total = max(float(vec[1]), 1.0)
            tp, fp, fn = vec[2:2 + ncls], vec[2 + ncls:2 + 2 * ncls], vec[2 + 2 * ncls:2 + 3 * ncls]
            stats[f"{split}_acc"] = float(vec[0]) / total
            stats[f"{split}_loss"] = float(vec[-1]) / total
            stats[f"{split}_precision"] = float((tp / (tp + fp).clamp_min(1)).mean())
            stats[f"{split}_recall"] = float((tp / (tp + fn).clamp_min(1)).mean())
        stats["eval_path"] = "native"
        self.history.append(stats)
        if self.metrics is not None:
            self.metrics.log({"Train/Acc": stats["train_acc"], "Train/Loss": stats["train_loss"],
                              "Test/Acc": stats["test_acc"], "Test/Loss": stats["test_loss"], "round": round_idx})
        return stats

    # ------------------------------------------------------------------ driver
    def train(self):
        args = self.args
        start = 0
        if bool(getattr(args, "resume", False)) and getattr(args, "checkpoint_dir", None):
            start = self.resume(args.checkpoint_dir)
        for round_idx in range(start, int(args.comm_round)):
            self.train_one_round(round_idx)
            freq = int(getattr(args, "frequency_of_the_test", 1) or 1)
            if freq > 0 and (round_idx == int(args.comm_round) - 1 or round_idx % freq == 0):
                self.evaluate(round_idx)
            every = int(getattr(args, "checkpoint_every", 0) or 0)
            if every and getattr(args, "checkpoint_dir", None) and (round_idx + 1) % every == 0:
                self.save(args.checkpoint_dir)          # collective: sharded optimizer state is gathered from every rank
        return self.global_state_dict()

    def global_state_dict(self):
        return self.layout.unflatten(self.global_flat.detach().cpu())

    def save(self, cdir):
        from ...utils.checkpoint import save_run_state
        extra = {"engine": self.engine.state_dict()} if self.engine is not None else \
            {"server_opt": self.server_opt.state_dict() if self.server_opt is not None else None}
        if self.rank == 0:
            save_run_state(cdir, self.round_idx, self.global_state_dict(), extra)

    def resume(self, cdir) -> int:
        from ...utils.checkpoint import load_run_state
        st = load_run_state(cdir, restore_rng=False)
        if st is None:
            return 0
        flat = self.layout.flatten(st["state_dict"])
        if self.engine is not None:
            self.engine.set_global(flat)
            if st["extra"].get("engine"):
                self.engine.load_state_dict(st["extra"]["engine"])
        else:
            self.global_flat.copy_(flat)
            if self.server_opt is not None and st["extra"].get("server_opt"):
                self.server_opt.load_state_dict(st["extra"]["server_opt"])
        self.round_idx = int(st["round"])
        return self.round_idx
