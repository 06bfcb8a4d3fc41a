"""Pairwise-mask secure aggregation (Bonawitz et al. 2017 style) on quantised updates.

The reference only ships the finite-field toolbox; its TurboAggregate protocol body is ``pass``
(TA_trainer.py:106-116).  Protocol implemented here (honest-but-curious server, dropout
recovery through Shamir shares of each client's mask seed):

  1. every pair (i, j), i < j, agrees on a seed s_ij (Diffie-Hellman over GF(p) with generator g);
  2. client i uploads  y_i = q(n_i/N * x_i) + sum_{j>i} PRG(s_ij) - sum_{j<i} PRG(s_ji)   (mod p);
  3. the server adds the y_i: masks cancel pairwise and the sum de-quantises to the weighted mean;
  4. for a dropped client d the survivors reveal their seeds with d and the server removes the orphan masks.

Mask expansion + modular adds run in ``fb_ff_mask`` / ``fb_ff_add`` (csrc/secagg.cu) on CUDA tensors (Philox
counter PRG, masks never materialised in memory) and in NumPy otherwise — both use the same PRG definition
so shares interoperate."""
from __future__ import annotations

from typing import Dict, List, Sequence

import numpy as np
import torch

from .mpc_function import DEFAULT_PRIME, my_key_agreement, my_pk_gen

_M0, _M1, _W0, _W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85


def _philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10 (NumPy uint64 lanes), bit-identical to csrc/common.cuh."""
    mask = np.uint64(0xFFFFFFFF)
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) for c in (c0, c1, c2, c3))
    k0, k1 = np.uint64(k0), np.uint64(k1)
    for _ in range(10):
        p0, p1 = np.uint64(_M0) * c0, np.uint64(_M1) * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & mask, p1 >> np.uint64(32), p1 & mask
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & mask, lo1, (hi0 ^ c3 ^ k1) & mask, lo0
        k0, k1 = (k0 + np.uint64(_W0)) & mask, (k1 + np.uint64(_W1)) & mask
    return c0, c1, c2, c3


def prg_mask(seed: int, n: int, p: int) -> np.ndarray:
    """n field elements from seed: word i = Philox(counter=i//4, stream 0x5ec0a66, key=seed)[i%4] mod p."""
    nv = (n + 3) // 4
    v = np.arange(nv, dtype=np.uint64)
    out = _philox4x32_10(v & np.uint64(0xFFFFFFFF), v >> np.uint64(32), np.full(nv, 0x5EC0A66, dtype=np.uint64),
                         np.zeros(nv, dtype=np.uint64), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    return (np.stack(out, axis=1).reshape(-1)[:n] % np.uint64(p)).astype(np.int64)


class SecAggClient:
    def __init__(self, cid: int, p: int = DEFAULT_PRIME, g: int = 7, rng=None):
        self.cid, self.p, self.g = cid, p, g
        rng = rng or np.random.default_rng(cid + 12345)
        self.sk = int(rng.integers(2, p - 2))
        self.pk = my_pk_gen(self.sk, p, g)
        self.seeds: Dict[int, int] = {}

    def agree(self, public_keys: Dict[int, int]):
        for other, pk in public_keys.items():
            if other != self.cid:
                self.seeds[other] = my_key_agreement(self.sk, pk, self.p, self.g)

    def mask(self, x: torch.Tensor, weight: float, scale: float, participants: Sequence[int]) -> torch.Tensor:
        """Quantise ``weight * x`` and add the pairwise masks towards every other participant."""
        others = [o for o in participants if o != self.cid]
        seeds = [self.seeds[o] for o in others]
        signs = [1 if o > self.cid else -1 for o in others]
        if x.is_cuda:
            from ... import ops
            q = ops.ff_quantize((x * weight).contiguous(), scale, self.p)
            return ops.ff_mask(q, seeds, signs, self.p)
        q = np.mod(np.rint((x.double() * weight).numpy() * scale).astype(np.int64), self.p)
        for s, sg in zip(seeds, signs):
            q = np.mod(q + sg * prg_mask(s, q.size, self.p).reshape(q.shape), self.p)
        return torch.from_numpy(q.astype(np.int64))

    def reveal_seed(self, dropped: int) -> int:
        return self.seeds[dropped]


class SecAggServer:
    def __init__(self, p: int = DEFAULT_PRIME):
        self.p = p

    def aggregate(self, masked: Dict[int, torch.Tensor], scale: float, participants: Sequence[int],
                  revealed: Dict[int, Dict[int, int]] = None) -> torch.Tensor:
        """Sum the uploads of the survivors; ``revealed[d][i]`` = seed between dropped client d and survivor i."""
        ids = sorted(masked)
        first = masked[ids[0]]
        if first.is_cuda:
            from ... import ops
            acc = first.clone()
            for i in ids[1:]:
                ops.ff_add_(acc, masked[i], self.p)
            for d, seeds in (revealed or {}).items():
                # survivor i added sign(d - i) * PRG(s_id); remove it
                s, sg = list(seeds.values()), [-(1 if d > i else -1) for i in seeds]
                acc = ops.ff_mask(acc, s, sg, self.p)
            return ops.ff_dequantize(acc, scale, self.p)
        acc = np.zeros(first.numel(), dtype=np.int64)
        for i in ids:
            acc = np.mod(acc + masked[i].numpy().reshape(-1), self.p)
        for d, seeds in (revealed or {}).items():
            for i, s in seeds.items():
                acc = np.mod(acc - (1 if d > i else -1) * prg_mask(s, acc.size, self.p), self.p)
        signed = np.where(acc > self.p // 2, acc - self.p, acc)
        return torch.from_numpy(signed.astype(np.float64) / scale).to(torch.float32).view(first.shape)


def secure_weighted_mean(updates: List[torch.Tensor], weights: Sequence[float], scale: float = 2 ** 16,
                         p: int = DEFAULT_PRIME, dropped: Sequence[int] = ()) -> torch.Tensor:
    """End-to-end helper (and test oracle): masked uploads -> server sum -> weighted mean of the survivors."""
    n = len(updates)
    clients = [SecAggClient(i, p) for i in range(n)]
    pks = {c.cid: c.pk for c in clients}
    for c in clients:
        c.agree(pks)
    parts = list(range(n))
    alive = [i for i in parts if i not in set(dropped)]
    total = float(sum(weights[i] for i in alive))
    masked = {i: clients[i].mask(updates[i], weights[i] / total, scale, parts) for i in alive}
    revealed = {d: {i: clients[i].reveal_seed(d) for i in alive} for d in dropped}
    return SecAggServer(p).aggregate(masked, scale, parts, revealed)
