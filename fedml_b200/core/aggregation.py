"""Plain-PyTorch aggregation maths on ``state_dict``s.

These functions are (a) the product path on CPU / in the single-process simulator and
(b) the fp32/fp64 oracles every CUDA kernel in ``csrc/`` is tested against.  They restate
the numerics of the reference:

  * FedAvg  — ``w = sum_i (n_i / sum n) * w_i`` over every state entry incl. BN buffers and
    the int64 ``num_batches_tracked`` (fedavg/fedavg_api.py:206-221);
  * FedOpt  — pseudo-gradient ``g = w_global - w_avg`` on named parameters, buffers are
    overwritten by the average (fedopt/fedopt_api.py:173-186);
  * FedNova — ``w <- w - tau_eff * sum_i norm_grad_i`` with optional global momentum
    (fednova/fednova_trainer.py:136-164);
  * clipping / weak DP / coordinate median — core/robustness/robust_aggregation.py:47-99
    (with the two reference bugs of SURVEY §0.3 items 4-5 fixed);
  * trimmed mean (Yin et al. 2018) and (multi-)Krum (Blanchard et al. 2017), which the
    reference lacks.
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Dict, Iterable, List, Sequence, Tuple

import torch

StateDict = Dict[str, torch.Tensor]


def is_weight_param(name: str) -> bool:
    return ("running_mean" not in name and "running_var" not in name
            and "num_batches_tracked" not in name)


def vectorize_weight(state_dict: StateDict, weights_only: bool = True) -> torch.Tensor:
    parts = [v.reshape(-1).to(torch.float32) for k, v in state_dict.items()
             if (is_weight_param(k) or not weights_only)]
    return torch.cat(parts) if parts else torch.zeros(0)


def unvectorize_like(vec: torch.Tensor, like: StateDict, weights_only: bool = True) -> StateDict:
    out, off = OrderedDict(), 0
    for k, v in like.items():
        if is_weight_param(k) or not weights_only:
            n = v.numel()
            out[k] = vec[off:off + n].view(v.shape).to(v.dtype if v.dtype.is_floating_point else torch.float32)
            off += n
        else:
            out[k] = v.clone()
    return out


# --------------------------------------------------------------------------- FedAvg
def fedavg_aggregate(w_locals: Sequence[Tuple[float, StateDict]], total: float = None) -> StateDict:
    """Sample-count weighted mean.  ``total`` overrides the denominator (FedNAS and
    TurboAggregate-SP divide by the global train count instead)."""
    denom = float(sum(n for n, _ in w_locals)) if total is None else float(total)
    first = w_locals[0][1]
    out = OrderedDict()
    for k, ref in first.items():
        acc = None
        for n, sd in w_locals:
            term = sd[k].to(torch.float32) * (float(n) / denom) if not sd[k].dtype.is_floating_point \
                else sd[k] * (float(n) / denom)
            acc = term.clone() if acc is None else acc.add_(term)
        out[k] = acc
    return out


def uniform_aggregate(w_locals: Sequence[Tuple[float, StateDict]]) -> StateDict:
    """Unweighted mean (``_aggregate_noniid_avg``, fedavg_api.py:223-236)."""
    return fedavg_aggregate([(1.0, sd) for _, sd in w_locals])


def cast_like(agg: StateDict, like: StateDict) -> StateDict:
    """Cast an aggregate back to the dtypes of a model state (int64 counters etc.)."""
    return OrderedDict((k, v.to(like[k].dtype) if k in like else v) for k, v in agg.items())


# --------------------------------------------------------------------------- FedOpt
class ServerOptimizer:
    """FedOpt server step over a state_dict (SGD(+momentum) / Adam / Adagrad / Yogi).

    The update rules are those of ``torch.optim`` (which the reference instantiates by
    lower-cased class name, fedopt/optrepo.py:10-38) plus Yogi from Reddi et al. 2021.  State
    persists across rounds, which is what the reference emulates by saving/reloading the
    optimizer ``state_dict`` every round (fedopt_api.py:134-143).
    """

    def __init__(self, name="sgd", lr=1.0, momentum=0.0, beta1=0.9, beta2=0.99, eps=1e-3,
                 weight_decay=0.0, nesterov=False):
        self.name = str(name).lower()
        if self.name in ("fedadam",):
            self.name = "adam"
        if self.name in ("fedyogi",):
            self.name = "yogi"
        if self.name in ("fedadagrad",):
            self.name = "adagrad"
        if self.name not in ("sgd", "adam", "adagrad", "yogi"):
            raise ValueError(f"unsupported server optimizer {name!r}")
        self.lr, self.momentum, self.beta1, self.beta2, self.eps = lr, momentum, beta1, beta2, eps
        self.weight_decay, self.nesterov = weight_decay, nesterov
        self.step_count = 0
        self.m: Dict[str, torch.Tensor] = {}
        self.v: Dict[str, torch.Tensor] = {}

    def step(self, w_global: StateDict, w_avg: StateDict, param_names: Iterable[str]) -> StateDict:
        self.step_count += 1
        t = self.step_count
        names = set(param_names)
        out = OrderedDict()
        for k, avg in w_avg.items():
            if k not in names:
                out[k] = avg.clone()
                continue
            w = w_global[k].to(torch.float32)
            g = w - avg.to(torch.float32)
            if self.weight_decay:
                g = g + self.weight_decay * w
            if self.name == "sgd":
                if self.momentum:
                    buf = self.m.get(k)
                    buf = g.clone() if buf is None else buf.mul(self.momentum).add(g)
                    self.m[k] = buf
                    g = g + self.momentum * buf if self.nesterov else buf
                out[k] = w - self.lr * g
            elif self.name == "adagrad":
                v = self.v.get(k, torch.zeros_like(w)) + g * g
                self.v[k] = v
                out[k] = w - self.lr * g / (v.sqrt() + self.eps)
            else:
                m = self.m.get(k, torch.zeros_like(w)) * self.beta1 + (1 - self.beta1) * g
                v_old = self.v.get(k, torch.zeros_like(w))
                g2 = g * g
                if self.name == "adam":
                    v = v_old * self.beta2 + (1 - self.beta2) * g2
                else:  # yogi
                    v = v_old - (1 - self.beta2) * g2 * torch.sign(v_old - g2)
                self.m[k], self.v[k] = m, v
                bc1 = 1 - self.beta1 ** t
                bc2 = 1 - self.beta2 ** t
                denom = (v.sqrt() / math.sqrt(bc2)) + self.eps
                out[k] = w - (self.lr / bc1) * m / denom
        return out

    def state_dict(self):
        return {"name": self.name, "step": self.step_count, "m": self.m, "v": self.v,
                "hyper": dict(lr=self.lr, momentum=self.momentum, beta1=self.beta1, beta2=self.beta2,
                              eps=self.eps, weight_decay=self.weight_decay, nesterov=self.nesterov)}

    def load_state_dict(self, sd):
        self.step_count, self.m, self.v = sd["step"], dict(sd["m"]), dict(sd["v"])


# --------------------------------------------------------------------------- FedNova
def fednova_aggregate(w_global: StateDict, norm_grads: Sequence[StateDict], tau_effs: Sequence[float],
                      gmf: float = 0.0, momentum_buf: StateDict = None, lr: float = 1.0):
    """``w <- w - tau_eff * sum_i norm_grad_i`` (+ global momentum ``gmf``).

    ``norm_grad_i`` already carries the client ratio ``n_i/N`` and the ``1/a_i`` normaliser;
    ``tau_eff = sum_i tau_i`` (fednova_trainer.py:136-164).  Returns (new_state, momentum_buf).
    """
    tau_eff = float(sum(tau_effs))
    cum = OrderedDict()
    for k in norm_grads[0]:
        acc = None
        for ng in norm_grads:
            term = ng[k].to(torch.float32) * tau_eff
            acc = term.clone() if acc is None else acc.add_(term)
        cum[k] = acc
    out = OrderedDict()
    new_buf = OrderedDict() if gmf else None
    for k, w in w_global.items():
        if k not in cum:
            out[k] = w.clone(); continue
        if gmf:
            prev = momentum_buf[k] if (momentum_buf is not None and k in momentum_buf) else None
            buf = cum[k] / lr if prev is None else prev * gmf + cum[k] / lr
            new_buf[k] = buf
            out[k] = (w.to(torch.float32) - lr * buf).to(w.dtype if w.dtype.is_floating_point else torch.float32)
        else:
            out[k] = (w.to(torch.float32) - cum[k]).to(w.dtype if w.dtype.is_floating_point else torch.float32)
    return out, new_buf


# --------------------------------------------------------------------------- robust
def norm_diff_clip(local: StateDict, global_: StateDict, norm_bound: float) -> StateDict:
    """``w_g + (w_i - w_g) / max(1, ||w_i - w_g||_2 / norm_bound)`` over weight entries only."""
    diff = vectorize_weight(local) - vectorize_weight(global_)
    scale = 1.0 / max(1.0, float(torch.linalg.vector_norm(diff)) / float(norm_bound))
    out, off = OrderedDict(), 0
    for k, v in local.items():
        if is_weight_param(k):
            n = v.numel()
            out[k] = (global_[k].to(torch.float32) + diff[off:off + n].view(v.shape) * scale).to(v.dtype)
            off += n
        else:
            out[k] = v.clone()
    return out


def add_gaussian_noise(local: StateDict, stddev: float, generator: torch.Generator = None) -> StateDict:
    out = OrderedDict()
    for k, v in local.items():
        if is_weight_param(k) and v.dtype.is_floating_point:
            out[k] = v + torch.randn(v.shape, generator=generator, device=v.device, dtype=v.dtype) * stddev
        else:
            out[k] = v.clone()
    return out


def _stack(w_locals, weights_only=False):
    return torch.stack([vectorize_weight(sd, weights_only=weights_only) for _, sd in w_locals], dim=0)


def coordinate_median(w_locals: Sequence[Tuple[float, StateDict]]) -> StateDict:
    """Coordinate-wise median over *all* entries (fixes SURVEY §0.3 item 5); for even K the
    lower of the two middle values, like ``torch.median`` (robust_aggregation.py:88-89)."""
    mat = _stack(w_locals)                       # [K, P]
    med = torch.median(mat, dim=0).values
    return unvectorize_like(med, w_locals[0][1], weights_only=False)


def trimmed_mean(w_locals: Sequence[Tuple[float, StateDict]], trim_ratio: float) -> StateDict:
    """Coordinate-wise beta-trimmed mean: drop the floor(beta*K) largest and smallest values of
    every coordinate, average the rest (Yin et al., ICML 2018)."""
    mat = _stack(w_locals)
    K = mat.shape[0]
    b = int(math.floor(trim_ratio * K))
    if 2 * b >= K:
        raise ValueError("trim_ratio removes every client")
    srt = torch.sort(mat, dim=0).values
    kept = srt[b:K - b].to(torch.float64).mean(dim=0).to(torch.float32)
    return unvectorize_like(kept, w_locals[0][1], weights_only=False)


def krum_scores(mat: torch.Tensor, f: int) -> torch.Tensor:
    """Krum score of each row: sum of squared distances to its K-f-2 nearest other rows."""
    K = mat.shape[0]
    m64 = mat.to(torch.float64)
    sq = (m64 * m64).sum(dim=1)
    d2 = (sq[:, None] + sq[None, :] - 2.0 * (m64 @ m64.t())).clamp_min_(0.0)
    d2.fill_diagonal_(float("inf"))
    nn = max(1, K - f - 2)
    return torch.sort(d2, dim=1).values[:, :nn].sum(dim=1)


def krum_select(scores: torch.Tensor, m: int = 1) -> List[int]:
    return torch.argsort(scores, stable=True)[:m].tolist()


def krum(w_locals: Sequence[Tuple[float, StateDict]], f: int, m: int = 1) -> StateDict:
    """(Multi-)Krum (Blanchard et al., NeurIPS 2017): pick the m clients with the smallest
    scores, then average them (sample-weighted for m > 1, the single winner for m = 1)."""
    mat = _stack(w_locals, weights_only=True)
    chosen = krum_select(krum_scores(mat, f), m)
    return fedavg_aggregate([w_locals[i] for i in chosen])


# --------------------------------------------------------------------------- byzantine injection
def apply_byzantine(local: StateDict, global_: StateDict, kind: str, scale: float = 10.0,
                    generator: torch.Generator = None) -> StateDict:
    """Fault-injection harness used to exercise the robust aggregators."""
    out = OrderedDict()
    for k, v in local.items():
        if not (is_weight_param(k) and v.dtype.is_floating_point):
            out[k] = v.clone(); continue
        g = global_[k].to(v.dtype)
        if kind == "sign_flip":
            out[k] = g - (v - g)
        elif kind == "scaled":
            out[k] = g + scale * (v - g)
        elif kind == "gaussian":
            out[k] = g + torch.randn(v.shape, generator=generator, dtype=v.dtype) * scale
        elif kind == "zero":
            out[k] = torch.zeros_like(v)
        else:
            raise ValueError(f"unknown byzantine attack {kind!r}")
    return out
