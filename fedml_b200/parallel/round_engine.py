"""Driver of the fused FL round kernel (csrc/round.cu).

One ``RoundEngine`` per rank owns the peer-visible buffers (global model, partial sums, flags),
the server-optimizer moments and the DP scratch, and turns "these C client models with these
weights" into one cooperative kernel launch that leaves the new global model on every rank.
It replaces the reference's aggregate-on-rank-0-then-send path (FedAVGAggregator.aggregate +
MPI model messages) — there is no NCCL call on the round path.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional

import torch

from .. import ops
from ..ops.native import FB_OPT, FB_ROUND_MAX_BLOCKS, FbRoundParams
from .arena import ArenaLayout
from .symmetric import SymmetricRegion


@dataclass
class ServerOptConfig:
    name: str = "none"          # none | sgd | adam | adagrad | yogi
    lr: float = 1.0
    momentum: float = 0.0
    beta1: float = 0.9
    beta2: float = 0.99
    eps: float = 1e-3
    nesterov: bool = False

    @classmethod
    def from_args(cls, args, enabled: bool):
        if not enabled:
            return cls()
        name = str(getattr(args, "server_optimizer", "sgd")).lower()
        name = {"fedadam": "adam", "fedyogi": "yogi", "fedadagrad": "adagrad"}.get(name, name)
        return cls(name, float(getattr(args, "server_lr", 1.0)), float(getattr(args, "server_momentum", 0.0)),
                   float(getattr(args, "server_beta1", 0.9)), float(getattr(args, "server_beta2", 0.99)),
                   float(getattr(args, "server_eps", 1e-3)))


@dataclass
class DPConfig:
    clip: float = 0.0
    sigma: float = 0.0
    seed: int = 0
    aggregate_noise: bool = False   # one draw on the aggregate (same output distribution, no per-client RNG cost)


def shard_bounds(P: int, P_train: int, G: int, rank: int):
    """Element range of the server-optimizer state that ``rank`` owns in two-shot mode: shard ``rank`` of the float4
    lanes of the flat model (csrc/round.cu: S4 = ceil(P/4 / G) lanes per shard), clipped to the trainable prefix."""
    nvec = P // 4
    s4 = (nvec + G - 1) // G
    lo, hi = min(rank * s4 * 4, P_train), min((rank + 1) * s4 * 4, P_train)
    return lo, hi


def merge_sharded_moments(per_rank, P: int, P_train: int):
    """Two-shot mode updates the moments of shard r on rank r only: assemble the complete tensor from every rank's own
    shard (``per_rank``: list of [P_train] tensors, one per rank)."""
    G = len(per_rank)
    out = torch.zeros_like(per_rank[0])
    for r, t in enumerate(per_rank):
        lo, hi = shard_bounds(P, P_train, G, r)
        out[lo:hi] = t[lo:hi]
    return out


class RoundEngine:
    TWO_SHOT_BYTES = 8 << 20    # above this the reduce is bandwidth-bound: shard it

    def __init__(self, layout: ArenaLayout, device, group=None, opt: Optional[ServerOptConfig] = None,
                 dp: Optional[DPConfig] = None, two_shot: Optional[bool] = None, max_clients: int = 1024,
                 provider: str = "auto", use_multicast: bool = True):
        self.layout, self.device = layout, torch.device(device)
        self.P, self.P_train = layout.P, layout.P_train
        self.opt = opt or ServerOptConfig()
        self.dp = dp or DPConfig()
        flag_bytes = ops.round_flag_bytes(64)
        need = 2 * self.P * 4 + flag_bytes + 8 * 64 * 4 + 4096
        self.region = SymmetricRegion(need, self.device, group, provider)
        self.G, self.rank = self.region.world, self.region.rank
        self.region.carve("global", self.P * 4)
        self.region.carve("partial", self.P * 4)
        self.region.carve("flags", flag_bytes)
        self.region.carve("barflags", 8 * 64 * 4)
        self.global_model = self.region.tensor("global")
        self.partial = self.region.tensor("partial")
        self.region.tensor("flags", torch.int32).zero_()
        self.two_shot = (self.P * 4 >= self.TWO_SHOT_BYTES) if two_shot is None else bool(two_shot)
        if self.G > 1:
            self._peer_partial = self.region.peer_table("partial")
            self._peer_global = self.region.peer_table("global")
            self._peer_flags = self.region.peer_table("flags")
            self._peer_bar = self.region.peer_table("barflags")
            mc = self.region.multicast_ptr if use_multicast else 0
            self._mc_partial = mc + self.region._names["partial"][0] if mc else 0
            self._mc_global = mc + self.region._names["global"][0] if mc else 0
        else:
            self._peer_partial = self._peer_global = self._peer_flags = None
            self._mc_partial = self._mc_global = 0
        z = lambda: torch.zeros(self.P_train, dtype=torch.float32, device=self.device)
        self.m = z() if (self.opt.name in ("adam", "yogi") or (self.opt.name == "sgd" and self.opt.momentum)) else None
        self.v = z() if self.opt.name in ("adam", "yogi", "adagrad") else None
        self.grid = ops.round_grid(self.P, self.G)
        self.max_clients = max_clients
        self.sqnorm_part = torch.zeros(self.grid * max_clients, dtype=torch.float32, device=self.device)
        self.clip_scale = torch.zeros(max_clients, dtype=torch.float32, device=self.device)
        self.grid_bar = torch.zeros(4, dtype=torch.int32, device=self.device)
        self.epoch = 0
        self.bar_epoch = 0
        self.opt_step = 0
        self.clip_calls = 0
        self.region.barrier()

    # ------------------------------------------------------------------
    def device_barrier(self, slot: int = 0):
        """Cross-rank barrier executed on the stream (flag writes over peer memory, no host sync)."""
        if self.G > 1:
            self.bar_epoch += 1
            ops.peer_barrier(self._peer_bar, self.rank, self.G, self.bar_epoch, slot)

    def set_global(self, flat: torch.Tensor):
        self.global_model.copy_(flat.to(self.device, torch.float32))

    def aggregate(self, clients: Optional[torch.Tensor], weights: Optional[torch.Tensor], inv_total: float,
                  client_base: int = 0, round_idx: int = 0, weights_buf: Optional[torch.Tensor] = None,
                  inv_total_buf: Optional[float] = None, self_coef: float = 0.0, grad_scale: float = 1.0,
                  use_opt: bool = True, weight_sq_sum: Optional[float] = None) -> torch.Tensor:
        """Launch the fused round.  ``clients``: fp32 [C, ld>=P] rows (or None when this rank drew no
        client this round); ``weights``: fp32 [C] on the device.  Returns the (updated in place)
        global model of this rank."""
        p = FbRoundParams()
        Cn = 0 if clients is None else clients.shape[0]
        if Cn > self.max_clients:
            raise ValueError(f"{Cn} clients exceed max_clients={self.max_clients}")
        if Cn:
            p.clients = clients.data_ptr()
            p.ld = clients.stride(0) if Cn > 1 else max(clients.stride(0), self.P)
            p.weights = weights.data_ptr()
            p.weights_buf = weights_buf.data_ptr() if weights_buf is not None else None
        p.C, p.client_base = Cn, client_base
        p.P, p.P_train = self.P, self.P_train
        p.inv_total = inv_total
        p.inv_total_buf = inv_total if inv_total_buf is None else inv_total_buf
        p.self_coef = self_coef
        p.global_model, p.partial = self.global_model.data_ptr(), self.partial.data_ptr()
        if self.G > 1:
            p.peer_partial, p.peer_global = self._peer_partial.data_ptr(), self._peer_global.data_ptr()
            p.peer_flags = self._peer_flags.data_ptr()
        p.rank, p.G = self.rank, self.G
        self.epoch += 1
        p.epoch, p.two_shot = self.epoch, int(self.two_shot and self.G > 1)
        p.clip_bound, p.noise_std = self.dp.clip, self.dp.sigma
        if self.dp.aggregate_noise and self.dp.sigma > 0 and weight_sq_sum is not None:
            # sum_i (w_i * inv_total) * sigma * z_i  ~  N(0, sigma^2 * inv_total^2 * sum_i w_i^2)
            p.noise_std, p.agg_noise_std = 0.0, self.dp.sigma * inv_total * float(weight_sq_sum) ** 0.5
        p.seed, p.round = self.dp.seed, round_idx
        if p.two_shot and self._mc_partial:
            p.mc_partial, p.mc_global = self._mc_partial, self._mc_global
        o = self.opt if use_opt else ServerOptConfig()
        p.opt = FB_OPT[o.name]
        if p.opt:
            self.opt_step += 1
        p.nesterov, p.lr, p.momentum = int(o.nesterov), o.lr, o.momentum
        p.beta1, p.beta2, p.eps = o.beta1, o.beta2, o.eps
        p.bias_corr1 = 1.0 - o.beta1 ** max(self.opt_step, 1)
        p.bias_corr2 = 1.0 - o.beta2 ** max(self.opt_step, 1)
        p.grad_scale = grad_scale
        p.m = self.m.data_ptr() if self.m is not None else None
        p.v = self.v.data_ptr() if self.v is not None else None
        if self.dp.clip > 0:
            # the kernel runs its intra-rank grid barrier (one arrival per block) only when it has client rows to clip:
            # count exactly those launches, or a round without clients leaves the target ahead of the counter and the
            # next clipped round spins forever
            if Cn > 0:
                self.clip_calls += 1
            p.sqnorm_part, p.clip_scale_out = self.sqnorm_part.data_ptr(), self.clip_scale.data_ptr()
            p.grid_bar, p.grid_bar_target = self.grid_bar.data_ptr(), self.grid * self.clip_calls
        ops.round_launch(p, self.device)
        return self.global_model

    # ------------------------------------------------------------------ failure detection
    @staticmethod
    def set_spin_timeout(ms: int):
        """Upper bound on any in-kernel wait for another rank (default 20 s)."""
        from ..ops.native import check, lib
        check(lib().fb_round_set_spin_timeout_ms(C.c_int64(int(ms))), "fb_round_set_spin_timeout_ms")

    @staticmethod
    def check_spin_errors(clear: bool = True):
        """Synchronises with the device and raises ``NativeError`` (FB_ERR_TIMEOUT) if a cross-rank wait of a round
        kernel or barrier gave up since the last check: a peer died or never launched.  The global model produced by
        that round is garbage and must be discarded (resume from the last checkpoint)."""
        from ..ops.native import check, lib
        out = (C.c_uint32 * 4)()
        check(lib().fb_round_spin_status(out, C.c_int32(int(clear))), "round kernel")

    # ------------------------------------------------------------------ checkpoint helpers
    def _full_moment(self, t: Optional[torch.Tensor]):
        """Complete server-optimizer moment tensor.  In two-shot mode with several GPUs each rank only ever updates its
        own shard (round.cu phase 3), so the other shards of its copy are stale: take every rank's own shard (masked
        all-reduce over the bootstrap process group).  COLLECTIVE in that mode: call on every rank."""
        if t is None:
            return None
        if not (self.G > 1 and self.two_shot):
            return t.detach().cpu()
        import torch.distributed as dist
        if not dist.is_initialized():
            raise RuntimeError("RoundEngine.state_dict: two-shot mode on several GPUs needs torch.distributed to gather "
                               "the sharded optimizer state")
        lo, hi = shard_bounds(self.P, self.P_train, self.G, self.rank)
        full = torch.zeros_like(t)
        full[lo:hi] = t[lo:hi]
        dist.all_reduce(full)
        return full.cpu()

    def state_dict(self):
        """Checkpoint state.  With G > 1 in two-shot mode this is a collective (see ``_full_moment``)."""
        return {"global": self.global_model.detach().cpu(), "m": self._full_moment(self.m), "v": self._full_moment(self.v),
                "opt_step": self.opt_step}

    def load_state_dict(self, sd):
        self.global_model.copy_(sd["global"])
        if self.m is not None and sd.get("m") is not None:
            self.m.copy_(sd["m"])
        if self.v is not None and sd.get("v") is not None:
            self.v.copy_(sd["v"])
        self.opt_step = int(sd.get("opt_step", 0))


def emulated_round(clients_per_rank, weights_per_rank, global_model: torch.Tensor, P_train: int, inv_total: float,
                   two_shot: bool, opt: ServerOptConfig = None, dp: DPConfig = None, m=None, v=None, opt_step=1,
                   round_idx=0, epoch=1, state=None, self_coef=0.0, grad_scale=1.0, weights_buf_per_rank=None,
                   inv_total_buf=None, dead_rank=None):
    """Run G emulated ranks of the round kernel in ONE cooperative launch on one GPU.

    Returns the list of per-rank global models after the round.  Used by the GPU tests to exercise
    the multi-rank protocol (flags, one-/two-shot, determinism) without several GPUs."""
    dev = global_model.device
    G = len(clients_per_rank)
    P = global_model.numel()
    opt, dp = opt or ServerOptConfig(), dp or DPConfig()
    st = state if state is not None else {}
    if "globals" not in st:
        st["globals"] = [global_model.clone() for _ in range(G)]
        st["partials"] = [torch.zeros(P, device=dev) for _ in range(G)]
        st["flags"] = [torch.zeros(ops.round_flag_bytes(G) // 4, dtype=torch.int32, device=dev) for _ in range(G)]
        need_m = opt.name in ("adam", "yogi") or (opt.name == "sgd" and opt.momentum)
        st["m"] = [(m.clone() if m is not None else torch.zeros(P_train, device=dev)) if need_m else None for _ in range(G)]
        st["v"] = [(v.clone() if v is not None else torch.zeros(P_train, device=dev))
                   if opt.name in ("adam", "yogi", "adagrad") else None for _ in range(G)]
        st["grid_bar"] = [torch.zeros(4, dtype=torch.int32, device=dev) for _ in range(G)]
        st["clip_calls"] = [0] * G        # per rank: a rank without clients skips the grid barrier (see RoundEngine.aggregate)
    grid = ops.round_grid(P, G, G)
    tab = lambda ts: torch.tensor([t.data_ptr() for t in ts], dtype=torch.int64).to(dev)
    t_partial, t_global, t_flags = tab(st["partials"]), tab(st["globals"]), tab(st["flags"])
    params, keep = [], []
    base = 0
    for r in range(G):
        p = FbRoundParams()
        cl, wt = clients_per_rank[r], weights_per_rank[r]
        Cn = 0 if cl is None else cl.shape[0]
        if Cn:
            p.clients, p.ld, p.weights = cl.data_ptr(), (cl.stride(0) if Cn > 1 else max(cl.stride(0), P)), wt.data_ptr()
            if weights_buf_per_rank is not None:
                p.weights_buf = weights_buf_per_rank[r].data_ptr()
        p.C, p.client_base, p.P, p.P_train = Cn, base, P, P_train
        base += Cn
        p.inv_total = inv_total
        p.inv_total_buf = inv_total if inv_total_buf is None else inv_total_buf
        p.self_coef, p.grad_scale = self_coef, grad_scale
        p.global_model, p.partial = st["globals"][r].data_ptr(), st["partials"][r].data_ptr()
        p.peer_partial, p.peer_global, p.peer_flags = t_partial.data_ptr(), t_global.data_ptr(), t_flags.data_ptr()
        p.rank, p.G, p.epoch, p.two_shot = r, G, epoch, int(two_shot and G > 1)
        if dead_rank is not None and r == dead_rank:
            p.epoch = epoch - 1       # fault injection: this rank never announces the current round (a crashed peer)
        p.clip_bound, p.noise_std, p.seed, p.round = dp.clip, dp.sigma, dp.seed, round_idx
        p.opt, p.nesterov, p.lr, p.momentum = FB_OPT[opt.name], int(opt.nesterov), opt.lr, opt.momentum
        p.beta1, p.beta2, p.eps = opt.beta1, opt.beta2, opt.eps
        p.bias_corr1, p.bias_corr2 = 1.0 - opt.beta1 ** opt_step, 1.0 - opt.beta2 ** opt_step
        p.m = st["m"][r].data_ptr() if st["m"][r] is not None else None
        p.v = st["v"][r].data_ptr() if st["v"][r] is not None else None
        if dp.clip > 0:
            sq = torch.zeros(grid * max(Cn, 1), dtype=torch.float32, device=dev)
            cs = torch.zeros(max(Cn, 1), dtype=torch.float32, device=dev)
            keep += [sq, cs]
            p.sqnorm_part, p.clip_scale_out = sq.data_ptr(), cs.data_ptr()
            if Cn > 0:
                st["clip_calls"][r] += 1
            p.grid_bar, p.grid_bar_target = st["grid_bar"][r].data_ptr(), grid * st["clip_calls"][r]
        params.append(p)
    ops.round_launch_emulated(params, P, dev)
    st["_keep"] = keep
    return st["globals"], st
