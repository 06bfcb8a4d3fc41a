"""Byzantine-robust aggregation across GPUs (coordinate median, trimmed mean, Krum).

The K client models of a round stay where they were trained — rows of each rank's peer-visible
stack.  Median / trimmed mean: rank r owns coordinate shard r, its selection kernel (csrc/robust.cu)
loads the K values of every coordinate straight from the owning GPUs (the NVLink "transpose" happens
in the loads) and stores the result into every rank's global model.  Krum: each rank packs its
coordinate shard of all K rows (centred on the global model, bf16 hi/lo split), forms the partial
K x K Gram matrix on tensor cores, the partial Grams are summed by the round kernel over peer memory,
and the winners are averaged by the ordinary fused round.  Reference: coordinate_median_agg
(core/robustness/robust_aggregation.py:65-99); trimmed mean / Krum are absent there.
"""
from __future__ import annotations

import math

import torch

from .. import ops


class _FlatLayout:
    def __init__(self, n):
        self.P = self.P_train = (n + 3) // 4 * 4


def _row_table(sim, K, sizes):
    """Pointer (as mapped in this process) of every sampled client's row, ordered by sampled position."""
    from ..core.schedule import schedule_clients
    assign = schedule_clients(sizes, sim.world)
    ptrs = [0] * K
    for q, positions in enumerate(assign):
        base = sim.stack.data_ptr() if sim.stack_region is None else sim.stack_region.peer_ptr("stack", q)
        for i, p in enumerate(positions):
            ptrs[p] = base + i * sim.P * 4
    return ops.ptr_table(ptrs, sim.device), assign


def aggregate(sim, kind, K, trim, f, m, my_pos, sizes, round_idx):
    eng = sim.engine
    G, rank, P = sim.world, sim.rank, sim.P
    rows, assign = _row_table(sim, K, sizes)
    eng.device_barrier(0)                                   # every rank's rows are final
    shard = (P // 4 + G - 1) // G * 4
    j0, j1 = min(rank * shard, P), min((rank + 1) * shard, P)
    if kind in ("median", "trimmed_mean"):
        outs = eng._peer_global if G > 1 else ops.ptr_table([eng.global_model.data_ptr()], sim.device)
        L = ops.run_length_for(sim.C_max) if K > 128 else 0
        if L:
            # many clients: every rank sorts runs of its OWN rows in place (local HBM), then the shard owners merge-select
            # over all G * C_max rows.  Rows of a rank that drew fewer than C_max clients are filled with +inf: they sort
            # to the end of their run and K_valid = K keeps them out of the order statistics.
            n_mine = len(my_pos)
            if n_mine < sim.C_max:
                sim.stack[n_mine:].fill_(float("inf"))
            own = ops.ptr_table([sim.stack.data_ptr() + i * sim.P * 4 for i in range(sim.C_max)], sim.device)
            ops.robust_sort_runs(own, sim.C_max, L, 0, P)
            eng.device_barrier(2)                           # every rank's runs are sorted
            allrows = []
            for q in range(G):
                base = sim.stack.data_ptr() if sim.stack_region is None else sim.stack_region.peer_ptr("stack", q)
                allrows += [base + i * sim.P * 4 for i in range(sim.C_max)]
            ops.robust_select_runs(ops.ptr_table(allrows, sim.device), G * sim.C_max, K, L, j0, j1, kind, trim, outs, G)
        else:
            ops.robust_select(rows, K, j0, j1, kind, trim, outs, G)
        eng.device_barrier(1)                               # all shards have landed everywhere
        return
    # ---- Krum
    Pt = sim.P_train
    j1k = min(j1, Pt)
    Kp = K
    gram_local = torch.zeros(_FlatLayout(K * K).P, dtype=torch.float32, device=sim.device)
    if j1k > j0:
        ld = (j1k - j0 + 63) // 64 * 64
        hi, lo = ops.krum_pack(rows, K, eng.global_model, j0, j1k, ld, split=True)
        gram = ops.gram_bf16(hi, lo) if hasattr(ops, "gram_bf16") else _gram_torch(hi, lo)
        gram_local[:K * K].copy_(gram.reshape(-1))
    if G > 1:
        if getattr(sim, "_gram_engine", None) is None or sim._gram_engine.P != gram_local.numel():
            from .round_engine import RoundEngine
            sim._gram_engine = RoundEngine(_FlatLayout(K * K), sim.device, two_shot=False, max_clients=1,
                                           provider=eng.region.provider)
        ge = sim._gram_engine
        one = torch.ones(1, dtype=torch.float32, device=sim.device)
        gram_all = ge.aggregate(gram_local.view(1, -1), one, 1.0)[:K * K].view(K, K)
    else:
        gram_all = gram_local[:K * K].view(K, K)
    scores = ops.krum_scores(gram_all.contiguous(), f)
    chosen = torch.argsort(scores, stable=True)[:m].tolist()          # K floats to the host: the selection itself
    sim.d2h_bytes += 4 * K
    weights = torch.zeros(sim.C_max, dtype=torch.float32)
    tot = 0.0
    for i, p in enumerate(my_pos):
        if p in chosen:
            weights[i] = sizes[p]
    tot = float(sum(sizes[p] for p in chosen))
    sim.last_krum_choice = chosen
    eng.aggregate(sim.stack, weights.to(sim.device), 1.0 / tot, 0, round_idx, use_opt=False)


def _gram_torch(hi, lo):
    """cuBLAS fallback used until the tcgen05 Gram kernel is built (recorded by the caller)."""
    h = hi.float()
    g = h @ h.t()
    if lo is not None:
        hl = h @ lo.float().t()
        g = g + hl + hl.t()
    return g
