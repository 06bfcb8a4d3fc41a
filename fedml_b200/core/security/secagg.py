"""Pairwise-mask secure aggregation (Bonawitz et al. 2017 style) on quantised updates.

The reference only ships the finite-field toolbox; its TurboAggregate protocol body is ``pass``
(TA_trainer.py:106-116).  Protocol implemented here (double masking, honest-but-curious server):

  0. client i draws a private key sk_i from the OS entropy source and publishes pk_i = g^sk_i; public keys are the ONLY
     key material that leaves a client (the server relays them); every pair derives s_ij by Diffie-Hellman over GF(p);
  1. per round r client i draws a fresh self-mask seed b_i and Shamir-shares b_i and sk_i (t-of-n) among the
     participants;
  2. client i uploads  y_i = q(n_i/N * x_i) + PRG(b_i, r) + sum_{j>i} PRG(s_ij, r) - sum_{j<i} PRG(s_ji, r)  (mod p);
     the round index is part of the PRG counter, so no mask is ever reused (a reused mask would leak
     q(x_i(r)) - q(x_i(r')) from two uploads);
  3. unmasking: for every SURVIVOR the participants hand the server their shares of b_i (never of sk_i), for every
     DROPPED client their shares of sk_d (never of b_d); the server reconstructs, removes the self-masks of the
     survivors and the orphaned pairwise masks of the dropped clients, and de-quantises the sum.  A server that lies
     about who dropped gets sk_d but then not b_d, so the upload y_d stays masked.

Limits, stated plainly: GF(p) with p < 2^31 (the reference toolbox's field, mpc_function.py) gives ~31-bit keys — a
demonstration of the protocol, not production-strength cryptography; shares travel in the clear in the in-process
helpers (a deployment encrypts share_{i->j} under s_ij); no defence against a malicious (actively deviating) server.

Mask expansion + modular adds run in ``fb_ff_mask`` / ``fb_ff_add`` (csrc/secagg.cu) on CUDA tensors (Philox
counter PRG, masks never materialised in memory) and in NumPy otherwise — both use the same PRG definition
so shares interoperate."""
from __future__ import annotations

import secrets
from typing import Dict, List, Optional, Sequence

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


def prg_mask(seed: int, n: int, p: int, nonce: int = 0) -> np.ndarray:
    """n field elements from (seed, nonce): word i = Philox(counter=(i//4, tag 0x5ec0a66, nonce), key=seed)[i%4] mod p."""
    nv = (n + 3) // 4
    v = np.arange(nv, dtype=np.uint64)
    out = _philox4x32_10(v & np.uint64(0xFFFFFFFF), v >> np.uint64(32), np.full(nv, 0x5EC0A66, dtype=np.uint64),
                         np.full(nv, int(nonce) & 0xFFFFFFFF, dtype=np.uint64), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    return (np.stack(out, axis=1).reshape(-1)[:n] % np.uint64(p)).astype(np.int64)


def shamir_share(secret: int, ids: Sequence[int], t: int, p: int, rng=None) -> Dict[int, int]:
    """t-of-n Shamir shares of ``secret`` in GF(p) for the participants ``ids`` (evaluation points id + 1)."""
    draw = (lambda: secrets.randbelow(p)) if rng is None else (lambda: int(rng.integers(0, p)))
    coef = [int(secret) % p] + [draw() for _ in range(max(t, 1) - 1)]
    out = {}
    for i in ids:
        x, acc = i + 1, 0
        for c in reversed(coef):
            acc = (acc * x + c) % p
        out[i] = acc
    return out


def shamir_reconstruct(shares: Dict[int, int], p: int) -> int:
    """Lagrange interpolation at 0 from {participant id: share}."""
    xs = [i + 1 for i in shares]
    secret = 0
    for i, yi in shares.items():
        xi, num, den = i + 1, 1, 1
        for xj in xs:
            if xj != xi:
                num = (num * (-xj)) % p
                den = (den * (xi - xj)) % p
        secret = (secret + yi * num * pow(den, p - 2, p)) % p
    return secret


class SecAggClient:
    """One participant.  ``rng`` (a NumPy generator) makes the key material reproducible — for tests only; by default
    every secret comes from ``secrets`` (OS entropy) and nothing but the public key is derivable from the client id."""

    def __init__(self, cid: int, p: int = DEFAULT_PRIME, g: int = 7, rng=None):
        self.cid, self.p, self.g, self._rng = cid, p, g, rng
        self.sk = 2 + (secrets.randbelow(p - 4) if rng is None else int(rng.integers(0, p - 4)))
        self.pk = my_pk_gen(self.sk, p, g)
        self.seeds: Dict[int, int] = {}
        self.self_seed: Optional[int] = None
        self.held: Dict[int, Dict[str, int]] = {}      # shares other participants gave us: owner -> {"b": .., "sk": ..}

    def agree(self, public_keys: Dict[int, int]):
        for other, pk in public_keys.items():
            other = int(other)
            if other != self.cid:
                self.seeds[other] = my_key_agreement(self.sk, int(pk), self.p, self.g)

    # ---- double masking -------------------------------------------------------------------------------
    def new_round_secrets(self, participants: Sequence[int], t: Optional[int] = None):
        """Fresh self-mask seed for this round + Shamir shares of it and of sk for every participant:
        returns {peer: {"b": share, "sk": share}}."""
        n = len(participants)
        t = t or n // 2 + 1
        self.self_seed = secrets.randbelow(self.p) if self._rng is None else int(self._rng.integers(0, self.p))
        b = shamir_share(self.self_seed, participants, t, self.p, self._rng)
        sk = shamir_share(self.sk, participants, t, self.p, self._rng)
        return {j: {"b": b[j], "sk": sk[j]} for j in participants}

    def receive_shares(self, owner: int, shares: Dict[str, int]):
        self.held[int(owner)] = {"b": int(shares["b"]), "sk": int(shares["sk"])}

    def reveal_shares(self, survivors: Sequence[int], dropped: Sequence[int]) -> Dict[str, Dict[int, int]]:
        """What this participant tells the server at unmasking time: its share of b_i for survivors, of sk_d for dropped
        clients — never both for one client (refuses an id listed on both sides)."""
        both = set(survivors) & set(dropped)
        if both:
            raise ValueError(f"clients {sorted(both)} are claimed both alive and dropped: refusing to reveal shares")
        return {"b": {i: self.held[i]["b"] for i in survivors if i in self.held},
                "sk": {d: self.held[d]["sk"] for d in dropped if d in self.held}}

    def mask(self, x: torch.Tensor, weight: float, scale: float, participants: Sequence[int], round_idx: int = 0,
             self_mask: bool = False) -> torch.Tensor:
        """Quantise ``weight * x`` and add the pairwise masks towards every other participant (and, with ``self_mask``,
        the round's self mask PRG(b_i))."""
        others = [o for o in participants if o != self.cid]
        seeds = [self.seeds[o] for o in others]
        signs = [1 if o > self.cid else -1 for o in others]
        if self_mask:
            if self.self_seed is None:
                raise RuntimeError("mask(self_mask=True) needs new_round_secrets() first")
            seeds, signs = seeds + [self.self_seed], signs + [1]
        if x.is_cuda:
            from ... import ops
            q = ops.ff_quantize((x * weight).contiguous(), scale, self.p)
            return ops.ff_mask(q, seeds, signs, self.p, nonce=round_idx)
        q = np.mod(np.rint((x.double() * weight).numpy() * scale).astype(np.int64), self.p)
        for s, sg in zip(seeds, signs):
            q = np.mod(q + sg * prg_mask(s, q.size, self.p, round_idx).reshape(q.shape), self.p)
        return torch.from_numpy(q.astype(np.int64))

    def reveal_seed(self, dropped: int) -> int:
        """Single-masking shortcut (no self masks): hand over the pairwise seed with a dropped client."""
        return self.seeds[dropped]


class SecAggServer:
    def __init__(self, p: int = DEFAULT_PRIME, g: int = 7):
        self.p, self.g = p, g

    def aggregate(self, masked: Dict[int, torch.Tensor], scale: float, participants: Sequence[int],
                  revealed: Dict[int, Dict[int, int]] = None, round_idx: int = 0,
                  self_seeds: Optional[Dict[int, int]] = None) -> torch.Tensor:
        """Sum the uploads of the survivors.  ``revealed[d][i]`` = seed between dropped client d and survivor i;
        ``self_seeds[i]`` = reconstructed self-mask seed of survivor i (double masking)."""
        ids = sorted(masked)
        first = masked[ids[0]]
        # masks to take off: survivor i added sign(d - i) * PRG(s_id) for a dropped d, and +PRG(b_i)
        seeds, signs = [], []
        for d, sd in (revealed or {}).items():
            for i, s_ in sd.items():
                seeds.append(int(s_)); signs.append(-(1 if d > i else -1))
        for i, b in (self_seeds or {}).items():
            seeds.append(int(b)); signs.append(-1)
        if first.is_cuda:
            from ... import ops
            acc = first.clone()
            for i in ids[1:]:
                ops.ff_add_(acc, masked[i], self.p)
            if seeds:
                acc = ops.ff_mask(acc, seeds, signs, self.p, nonce=round_idx)
            return ops.ff_dequantize(acc, scale, self.p)
        acc = np.zeros(first.numel(), dtype=np.int64)
        for i in ids:
            acc = np.mod(acc + masked[i].numpy().reshape(-1), self.p)
        for s_, sg in zip(seeds, signs):
            acc = np.mod(acc + sg * prg_mask(s_, acc.size, self.p, round_idx), self.p)
        signed = np.where(acc > self.p // 2, acc - self.p, acc)
        return torch.from_numpy(signed.astype(np.float64) / scale).to(torch.float32).view(first.shape)

    def unmask_inputs(self, reveals: Dict[int, Dict[str, Dict[int, int]]], survivors: Sequence[int], dropped: Sequence[int],
                      public_keys: Dict[int, int], t: int):
        """From the participants' revealed shares: (self_seeds of survivors, pairwise seeds of dropped clients)."""
        self_seeds, pair = {}, {}
        for i in survivors:
            sh = {j: r["b"][i] for j, r in reveals.items() if i in r["b"]}
            if len(sh) < t:
                raise RuntimeError(f"only {len(sh)} of the {t} shares needed to remove the self mask of client {i}")
            self_seeds[i] = shamir_reconstruct(dict(list(sh.items())[:t]), self.p)
        for d in dropped:
            sh = {j: r["sk"][d] for j, r in reveals.items() if d in r["sk"]}
            if len(sh) < t:
                raise RuntimeError(f"only {len(sh)} of the {t} shares needed to recover the keys of dropped client {d}")
            sk_d = shamir_reconstruct(dict(list(sh.items())[:t]), self.p)
            pair[d] = {i: my_key_agreement(sk_d, int(public_keys[i]), self.p, self.g) for i in survivors}
        return self_seeds, pair


def secure_weighted_mean(updates: List[torch.Tensor], weights: Sequence[float], scale: float = 2 ** 16,
                         p: int = DEFAULT_PRIME, dropped: Sequence[int] = (), round_idx: int = 0, rng=None) -> torch.Tensor:
    """End-to-end helper (and test oracle): key agreement -> share distribution -> double-masked uploads -> share
    reveal -> server sum -> weighted mean of the survivors.  Runs all parties in this process."""
    n = len(updates)
    clients = [SecAggClient(i, p, rng=rng) for i in range(n)]
    pks = {c.cid: c.pk for c in clients}
    for c in clients:
        c.agree(pks)
    parts = list(range(n))
    t = n // 2 + 1
    for c in clients:                                   # share distribution (relayed by the server in a deployment)
        for j, sh in c.new_round_secrets(parts, t).items():
            clients[j].receive_shares(c.cid, sh)
    alive = [i for i in parts if i not in set(dropped)]
    total = float(sum(weights[i] for i in alive))
    masked = {i: clients[i].mask(updates[i], weights[i] / total, scale, parts, round_idx, self_mask=True) for i in alive}
    server = SecAggServer(p)
    reveals = {j: clients[j].reveal_shares(alive, list(dropped)) for j in alive}
    self_seeds, pair = server.unmask_inputs(reveals, alive, list(dropped), pks, t)
    return server.aggregate(masked, scale, parts, pair, round_idx, self_seeds)
