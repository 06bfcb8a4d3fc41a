"""S-FedAvg and HS-FedAvg: Shapley-value client valuation with biased sampling (the fork's additions;
reference: single_process/s_fedavg/fedavg_api.py:138-453, hs_fedavg/*.py).

Per round the Shapley value of every sampled client is computed with utility = accuracy (or macro
F1/recall/precision) of the aggregate of a client subset on a server validation set — exactly over all subsets
(``shapley_mode: exact``, :258-325) or by permutation Monte-Carlo until the estimate moves less than 0.005
(``isApproached``, :138-146, :196-255); ``phi_k <- alpha*phi_k + beta*SV_k`` (:255,320); next round's sampling
probability is proportional to ``exp(phi)`` when ``sampling_filter == "exp"`` (:435-453).  The client loss is
class-weighted cross-entropy (``calc_class_weight`` :112-136) with gradient clipping 1.0.  HS-FedAvg adds
amplitude normalisation of every batch and averages the running amplitude over clients each round."""
import itertools
import math
import time

import numpy as np
import torch

from ...core import aggregation as A
from .fedavg_api import FedAvgAPI
from .hs_fft import amplitude_normalize


class SFedAvgAPI(FedAvgAPI):
    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        n = int(self.args.client_num_in_total)
        self.phi = np.zeros(n)
        self.alpha, self.beta = float(getattr(self.args, "sv_alpha", 0.75)), float(getattr(self.args, "sv_beta", 0.25))
        self.mode = str(getattr(self.args, "shapley_mode", "mc"))
        self.metric = str(getattr(self.args, "sv_metric", "acc"))
        self.sv_history, self.sv_time = [], []
        self.args.client_grad_clip = float(getattr(self.args, "client_grad_clip", 1.0) or 1.0)
        self._val = self._make_validation()

    def _make_validation(self):
        batches = list(itertools.islice(iter(self.test_global), 8))
        return batches

    # ---- class-weighted client loss
    def calc_class_weight(self, train_data):
        """weight_c = n / (classes present on the client * samples of class c), 0 for absent classes
        (s_fedavg/fedavg_api.py:112-136): every class a client holds contributes equally to its loss."""
        ys = [b[1].reshape(-1) for b in train_data]
        ncls = int(getattr(self, "class_num", 0) or getattr(self.args, "class_num", 0) or 0)
        if not ys:
            return torch.zeros(max(ncls, 1))
        y = torch.cat(ys).to(torch.int64)
        ncls = max(ncls, int(y.max()) + 1)
        counts = torch.bincount(y, minlength=ncls).to(torch.float32)
        present = int((counts > 0).sum())
        w = torch.zeros(ncls)
        w[counts > 0] = float(y.numel()) / (present * counts[counts > 0])
        return w

    def _train_one_client(self, client, w_global):
        self.args.class_weight = self.calc_class_weight(client.local_training_data)
        try:
            return super()._train_one_client(client, w_global)
        finally:
            self.args.class_weight = None

    # ---- sampling biased by exp(phi)
    def _client_sampling(self, round_idx, n_total, n_round):
        if str(getattr(self.args, "sampling_filter", "exp")) != "exp" or round_idx == 0:
            return super()._client_sampling(round_idx, n_total, n_round)
        np.random.seed(round_idx)
        pr = np.exp(self.phi - self.phi.max()); pr /= pr.sum()
        return np.random.choice(range(n_total), min(n_round, n_total), replace=False, p=pr).tolist()

    # ---- utility of a subset
    def _utility(self, subset, w_locals, w_global):
        if not subset:
            sd = w_global
        else:
            sd = A.cast_like(A.fedavg_aggregate([w_locals[i] for i in subset]), w_global)
        self.model_trainer.set_model_params(sd)
        m = self.model_trainer.test(self._val, self.device, self.args)
        if self.metric == "acc":
            return m["test_correct"] / max(m["test_total"], 1)
        key = {"recall": "test_recall", "precision": "test_precision"}.get(self.metric, "test_recall")
        vals = list(m[key].values())
        return float(np.mean(vals)) if vals else 0.0

    def _shapley_exact(self, w_locals, w_global):
        K = len(w_locals)
        cache = {}
        util = lambda s: cache.setdefault(s, self._utility(list(s), w_locals, w_global))  # noqa: E731
        sv = np.zeros(K)
        for k in range(K):
            others = [i for i in range(K) if i != k]
            for r in range(K):
                for sub in itertools.combinations(others, r):
                    wgt = math.factorial(r) * math.factorial(K - r - 1) / math.factorial(K)
                    sv[k] += wgt * (util(tuple(sorted(sub + (k,)))) - util(tuple(sub)))
        return sv

    def _shapley_mc(self, w_locals, w_global, tol=0.005, max_perm=50):
        K = len(w_locals)
        rng = np.random.RandomState(len(self.sv_history))
        sv, prev = np.zeros(K), None
        for t in range(1, max_perm + 1):
            perm = rng.permutation(K)
            u_prev = self._utility([], w_locals, w_global)
            for j in range(K):
                u = self._utility(sorted(perm[:j + 1].tolist()), w_locals, w_global)
                sv[perm[j]] += (u - u_prev - sv[perm[j]]) / t if t > 1 else (u - u_prev)
                u_prev = u
            if prev is not None and np.linalg.norm(sv - prev) < tol:    # isApproached
                break
            prev = sv.copy()
        return sv

    def _aggregate(self, w_locals, w_global=None):
        t0 = time.time()
        sv = self._shapley_exact(w_locals, w_global) if self.mode == "exact" else self._shapley_mc(w_locals, w_global)
        self.sv_time.append(time.time() - t0)
        self.sv_history.append(sv)
        for slot, cid in enumerate(self._last_sampled):
            self.phi[cid] = self.alpha * self.phi[cid] + self.beta * sv[slot]
        return A.fedavg_aggregate(w_locals)

    def train(self):
        orig = self._client_sampling

        def remember(*a):
            self._last_sampled = orig(*a)
            return self._last_sampled

        self._client_sampling = remember
        return super().train()


class HSFedAvgAPI(SFedAvgAPI):
    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.mode = "exact" if int(self.args.client_num_per_round) <= 6 else self.mode
        self.amp_global = None

    def _train_one_client(self, client, w_global):
        data, amp = [], self.amp_global
        for x, y in client.local_training_data:
            if x.dim() == 4:
                # hs_fft_window: None (default) swaps the whole amplitude spectrum - the K14 kernel on the GPU; a float L applies
                # the reference's low-frequency window rule (its process() uses L = 0, i.e. only the DC term)
                x, amp, _ = amplitude_normalize(x, amp, window=getattr(self.args, "hs_fft_window", None))
            data.append((x, y))
        client.local_training_data = data
        self._client_amps.append(amp)
        return super()._train_one_client(client, w_global)

    def _aggregate(self, w_locals, w_global=None):
        amps = [a for a in self._client_amps if a is not None]
        if amps:
            self.amp_global = torch.stack(amps).mean(0)           # amplitude averaged over clients each round
        self._client_amps = []
        return super()._aggregate(w_locals, w_global)

    def train(self):
        self._client_amps = []
        return super().train()
