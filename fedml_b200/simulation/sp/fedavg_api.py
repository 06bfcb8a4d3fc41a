"""Single-process ("Parrot SP") simulators: FedAvg and the algorithms that share its skeleton.

Skeleton (reference: single_process/fedavg/fedavg_api.py:83-141): per round
``np.random.seed(round); np.random.choice(N, K, replace=False)`` -> every sampled client trains
from the global model -> aggregate -> periodic evaluation over *all* clients.  The reference has
nine near-identical copies of that loop; here one class owns the loop and each algorithm
overrides a hook:

  FedAvgAPI      - sample-weighted mean                       (fedavg_api.py:206-221)
  FedOptAPI      - + server optimizer on the pseudo-gradient  (fedopt/fedopt_api.py:134-186)
  FedProxAPI     - FedAvg with a real proximal local term     (absent in the reference)
  FedNovaAPI     - normalised averaging                       (fednova/fednova_trainer.py:136-164)
  FedAvgRobustAPI- clipping / weak-DP / median / trimmed mean / Krum, byzantine injection
  HierarchicalFLAPI - client -> group -> global               (hierarchical_fl/*.py)

Unlike the reference, parameters never leave the training device between client and server
steps, and the 8-item dataset tuple is accepted as is (the fork's 9-tuple with ``target_label``
is tolerated).
"""
from __future__ import annotations

import copy
import logging
import time
from collections import OrderedDict
from typing import Dict, List, Tuple

import numpy as np
import torch

from ...core import aggregation as A
from ...core.robustness import RobustAggregator
from ...utils.metrics import MetricsSink
from ..trainers import FedNovaOptimizer, create_model_trainer


def client_sampling(round_idx: int, client_num_in_total: int, client_num_per_round: int) -> List[int]:
    """Reference sampling rule, identical in all its copies (fedavg_api.py:179-193)."""
    if client_num_in_total == client_num_per_round:
        return list(range(client_num_in_total))
    num = min(client_num_per_round, client_num_in_total)
    np.random.seed(round_idx)
    return np.random.choice(range(client_num_in_total), num, replace=False).tolist()


class Client:
    """Reusable client slot (reference: fedavg/client.py:5-50)."""

    def __init__(self, client_idx, local_training_data, local_test_data, local_sample_number, args, device,
                 model_trainer):
        self.client_idx = client_idx
        self.local_training_data, self.local_test_data = local_training_data, local_test_data
        self.local_sample_number = local_sample_number
        self.args, self.device, self.model_trainer = args, device, model_trainer

    def update_local_dataset(self, client_idx, local_training_data, local_test_data, local_sample_number):
        self.client_idx = client_idx
        self.local_training_data, self.local_test_data = local_training_data, local_test_data
        self.local_sample_number = local_sample_number
        self.model_trainer.set_id(client_idx)

    def get_sample_number(self):
        return self.local_sample_number

    def train(self, w_global):
        self.model_trainer.set_model_params(w_global)
        self.model_trainer.train(self.local_training_data, self.device, self.args)
        return self.model_trainer.get_model_params()

    def local_test(self, b_use_test_dataset):
        data = self.local_test_data if b_use_test_dataset else self.local_training_data
        return self.model_trainer.test(data, self.device, self.args)


class FedAvgAPI:
    def __init__(self, args, device, dataset, model, model_trainer=None, server_aggregator=None):
        self.args, self.device = args, device
        ds = list(dataset)
        (self.train_data_num_in_total, self.test_data_num_in_total, self.train_global, self.test_global,
         self.train_data_local_num_dict, self.train_data_local_dict, self.test_data_local_dict,
         self.class_num) = ds[:8]
        self.extra_dataset_fields = ds[8:]
        self.val_global = None
        self.model = model
        self.model_trainer = model_trainer or create_model_trainer(model, args)
        self.server_aggregator = server_aggregator
        self.metrics = MetricsSink(args)
        self.round_times: List[float] = []
        self.client_list: List[Client] = []
        self._setup_clients()
        self.history: List[dict] = []

    # ------------------------------------------------------------------ set-up
    def _setup_clients(self):
        for client_idx in range(int(self.args.client_num_per_round)):
            cid = client_idx if client_idx in self.train_data_local_dict else next(iter(self.train_data_local_dict.keys()))
            self.client_list.append(Client(cid, self.train_data_local_dict[cid], self.test_data_local_dict[cid],
                                           self.train_data_local_num_dict[cid], self.args, self.device,
                                           self.model_trainer))

    def _client_sampling(self, round_idx, client_num_in_total, client_num_per_round):
        idx = client_sampling(round_idx, client_num_in_total, client_num_per_round)
        logging.info("client_indexes = %s", idx)
        return idx

    # ------------------------------------------------------------------ hooks
    def _train_one_client(self, client: Client, w_global):
        return client.get_sample_number(), client.train(copy.deepcopy(w_global))

    def _aggregate(self, w_locals: List[Tuple[float, dict]], w_global=None):
        if self.server_aggregator is not None:
            w_locals = self.server_aggregator.on_before_aggregation(w_locals)
            return self.server_aggregator.on_after_aggregation(self.server_aggregator.aggregate(w_locals))
        return A.fedavg_aggregate(w_locals)

    def _server_update(self, w_global, w_agg, round_idx):
        return w_agg

    # ------------------------------------------------------------------ main loop
    def train(self):
        args = self.args
        w_global = self.model_trainer.get_model_params()
        start_round = int(getattr(self, "_resume_round", 0))
        for round_idx in range(start_round, int(args.comm_round)):
            t0 = time.perf_counter()
            logging.info("################Communication round : %d", round_idx)
            client_indexes = self._client_sampling(round_idx, int(args.client_num_in_total),
                                                   int(args.client_num_per_round))
            w_locals = []
            for slot, client in enumerate(self.client_list[:len(client_indexes)]):
                cid = client_indexes[slot]
                client.update_local_dataset(cid, self.train_data_local_dict[cid], self.test_data_local_dict[cid],
                                            self.train_data_local_num_dict[cid])
                w_locals.append(self._train_one_client(client, w_global))
            w_agg = self._aggregate(w_locals, w_global)
            w_global = A.cast_like(self._server_update(w_global, w_agg, round_idx), w_global)
            self.model_trainer.set_model_params(w_global)
            self.round_times.append(time.perf_counter() - t0)
            freq = int(getattr(args, "frequency_of_the_test", 1) or 1)
            if round_idx == int(args.comm_round) - 1 or round_idx % freq == 0:
                self._local_test_on_all_clients(round_idx)
            self._maybe_checkpoint(round_idx, w_global)
        return w_global

    # ------------------------------------------------------------------ eval
    def _local_test_on_all_clients(self, round_idx):
        """Evaluate the global model on every client's train and test split (fedavg_api.py:238-326).
        ``args.ci`` truncates to one client, as the reference's ``--ci 1`` does."""
        agg = {"train": [0.0, 0.0, 0.0], "test": [0.0, 0.0, 0.0]}
        client = self.client_list[0]
        n_total = int(self.args.client_num_in_total)
        cap = int(getattr(self.args, "eval_client_cap", 0) or 0)
        ids = range(n_total) if not cap else range(min(cap, n_total))
        for cid in ids:
            if cid not in self.test_data_local_dict or self.test_data_local_dict[cid] is None:
                continue
            client.update_local_dataset(cid, self.train_data_local_dict[cid], self.test_data_local_dict[cid],
                                        self.train_data_local_num_dict[cid])
            for split, use_test in (("train", False), ("test", True)):
                m = client.local_test(use_test)
                agg[split][0] += m["test_correct"]; agg[split][1] += m["test_loss"]; agg[split][2] += m["test_total"]
            if int(getattr(self.args, "ci", 0)) == 1:
                break
        stats = {"round": round_idx,
                 "train_acc": agg["train"][0] / max(agg["train"][2], 1), "train_loss": agg["train"][1] / max(agg["train"][2], 1),
                 "test_acc": agg["test"][0] / max(agg["test"][2], 1), "test_loss": agg["test"][1] / max(agg["test"][2], 1)}
        self.history.append(stats)
        self.metrics.log({"Train/Acc": stats["train_acc"], "Train/Loss": stats["train_loss"],
                          "Test/Acc": stats["test_acc"], "Test/Loss": stats["test_loss"], "round": round_idx})
        logging.info("%s", stats)
        return stats

    def _validate_global_model(self, data=None):
        """Accuracy + per-label recall of the global model on the global test set
        (the fork's addition, fedavg_api.py:143-177)."""
        return self.model_trainer.test(data if data is not None else self.test_global, self.device, self.args)

    # ------------------------------------------------------------------ checkpoint
    def _maybe_checkpoint(self, round_idx, w_global):
        every = int(getattr(self.args, "checkpoint_every", 0) or 0)
        cdir = getattr(self.args, "checkpoint_dir", None)
        if every and cdir and (round_idx + 1) % every == 0:
            from ...utils.checkpoint import save_run_state
            save_run_state(cdir, round_idx + 1, w_global, self._extra_state())

    def _extra_state(self):
        return {}

    def resume(self, cdir=None):
        from ...utils.checkpoint import load_run_state
        st = load_run_state(cdir or self.args.checkpoint_dir)
        if st is None:
            return False
        self.model_trainer.set_model_params(A.cast_like(st["state_dict"], self.model_trainer.get_model_params()))
        self._resume_round = st["round"]
        self._load_extra_state(st.get("extra", {}))
        return True

    def _load_extra_state(self, extra):
        pass


# ============================================================================ FedOpt
class FedOptAPI(FedAvgAPI):
    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        args = self.args
        self.server_opt = A.ServerOptimizer(
            getattr(args, "server_optimizer", "sgd"), lr=float(getattr(args, "server_lr", 1.0)),
            momentum=float(getattr(args, "server_momentum", 0.0)), beta1=float(getattr(args, "server_beta1", 0.9)),
            beta2=float(getattr(args, "server_beta2", 0.99)), eps=float(getattr(args, "server_eps", 1e-3)))
        self._param_names = [n for n, _ in self.model.named_parameters()]

    def _server_update(self, w_global, w_agg, round_idx):
        return self.server_opt.step(w_global, w_agg, self._param_names)

    def _extra_state(self):
        return {"server_opt": self.server_opt.state_dict()}

    def _load_extra_state(self, extra):
        if "server_opt" in extra:
            self.server_opt.load_state_dict(extra["server_opt"])


# ============================================================================ FedProx
class FedProxAPI(FedAvgAPI):
    """FedAvg whose clients minimise ``F_i(w) + mu/2 ||w - w_global||^2`` (Li et al. 2020)."""

    def __init__(self, args, *a, **kw):
        if not float(getattr(args, "fedprox_mu", 0.0) or 0.0):
            args.fedprox_mu = float(getattr(args, "mu", 0.0) or 0.1)
        super().__init__(args, *a, **kw)


# ============================================================================ FedNova
class FedNovaAPI(FedAvgAPI):
    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.global_momentum_buffer = None
        self._param_names = [n for n, _ in self.model.named_parameters()]

    def _train_one_client(self, client: Client, w_global):
        args, device = self.args, self.device
        model = self.model_trainer.model.to(device)
        self.model_trainer.set_model_params(copy.deepcopy(w_global))
        model.train()
        ratio = client.get_sample_number() / float(self._round_total)
        lr = float(getattr(args, "learning_rate", getattr(args, "lr", 0.03)))
        opt = FedNovaOptimizer([p for p in model.parameters() if p.requires_grad], lr=lr, ratio=ratio,
                               gmf=float(getattr(args, "gmf", 0.0)), mu=float(getattr(args, "mu", 0.0)),
                               momentum=float(getattr(args, "momentum", 0.0)),
                               dampening=float(getattr(args, "dampening", 0.0)),
                               weight_decay=float(getattr(args, "wd", 0.0) or 0.0),
                               nesterov=bool(getattr(args, "nesterov", False)))
        crit = torch.nn.CrossEntropyLoss().to(device)
        for _ in range(int(args.epochs)):
            for x, y in client.local_training_data:
                x, y = x.to(device), y.to(device)
                opt.zero_grad(set_to_none=True)
                crit(model(x), y).backward()
                opt.step()
        a_i = opt.local_normalizing_vec
        w_now = self.model_trainer.get_model_params()
        norm_grad = OrderedDict()
        for k in w_now:
            if k in self._param_names_set:
                norm_grad[k] = (w_global[k].to(torch.float32) - w_now[k].to(torch.float32)) * (ratio / a_i)
        tau = opt.local_steps * ratio if float(getattr(args, "mu", 0.0)) != 0 else a_i * ratio
        return (client.get_sample_number(), {"norm_grad": norm_grad, "tau_eff": tau, "state": w_now})

    def train(self):
        self._param_names_set = set(self._param_names)
        return super().train()

    def _client_sampling(self, round_idx, n_total, n_round):
        idx = super()._client_sampling(round_idx, n_total, n_round)
        self._round_total = sum(self.train_data_local_num_dict[c] for c in idx)
        return idx

    def _aggregate(self, w_locals, w_global=None):
        grads = [d["norm_grad"] for _, d in w_locals]
        taus = [d["tau_eff"] for _, d in w_locals]
        lr = float(getattr(self.args, "learning_rate", getattr(self.args, "lr", 0.03)))
        new, self.global_momentum_buffer = A.fednova_aggregate(
            w_global, grads, taus, gmf=float(getattr(self.args, "gmf", 0.0)),
            momentum_buf=self.global_momentum_buffer, lr=lr)
        # buffers (BN statistics): plain sample-weighted mean, they have no gradient
        buf_avg = A.fedavg_aggregate([(n, {k: v for k, v in d["state"].items() if k not in self._param_names_set})
                                      for n, d in w_locals]) if any(k not in self._param_names_set for k in w_global) else {}
        new.update(buf_avg)
        return new


# ============================================================================ robust FedAvg
class FedAvgRobustAPI(FedAvgAPI):
    """FedAvg with byzantine clients and a defence (norm clipping, weak DP, coordinate median,
    trimmed mean, Krum).  ``byzantine_fraction`` of each round's sample is replaced by an attack
    from :func:`fedml_b200.core.aggregation.apply_byzantine`."""

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.robust = RobustAggregator(self.args)
        if not self.robust.defense_type and getattr(self.args, "robust_agg", "none") != "none":
            self.robust.defense_type = self.args.robust_agg
        self._gen = torch.Generator().manual_seed(int(getattr(self.args, "random_seed", 0)) + 1234)

    def _aggregate(self, w_locals, w_global=None):
        frac = float(getattr(self.args, "byzantine_fraction", 0.0) or 0.0)
        n_bad = int(round(frac * len(w_locals)))
        if n_bad:
            kind = str(getattr(self.args, "byzantine_attack", "sign_flip"))
            w_locals = [(n, A.apply_byzantine({k: v.cpu() for k, v in sd.items()},
                                              {k: v.cpu() for k, v in w_global.items()}, kind, generator=self._gen))
                        if i < n_bad else (n, sd) for i, (n, sd) in enumerate(w_locals)]
            dev = next(iter(w_global.values())).device
            w_locals = [(n, {k: v.to(dev) for k, v in sd.items()}) for n, sd in w_locals]
        return self.robust.defend(w_locals, w_global, self._gen if self.device.type == "cpu" else None)


# ============================================================================ hierarchical
class HierarchicalFLAPI(FedAvgAPI):
    """Two-level FedAvg: sampled clients are split into ``group_num`` groups; each global round
    runs ``group_comm_round`` FedAvg rounds inside every group and then averages the group models
    weighted by group sample counts (hierarchical_fl/trainer.py:15-70, group.py:41-70).  With
    full participation, full-batch clients and one local epoch, (global, group) = (2,5) and (5,2)
    reproduce flat FedAvg over 10 rounds — the reference's CI oracle (ci/CI-script-fedavg.sh:50-58)."""

    def _assign_groups(self, client_indexes):
        g = int(getattr(self.args, "group_num", 2))
        method = getattr(self.args, "group_method", "random")
        if method == "random":
            rng = np.random.RandomState(int(getattr(self.args, "random_seed", 0)))
            gid = rng.randint(0, g, int(self.args.client_num_in_total))
        else:
            gid = np.arange(int(self.args.client_num_in_total)) % g
        groups: Dict[int, List[int]] = {}
        for c in client_indexes:
            groups.setdefault(int(gid[c]), []).append(c)
        return groups

    def train(self):
        args = self.args
        w_global = self.model_trainer.get_model_params()
        client = self.client_list[0]
        for g_round in range(int(getattr(args, "global_comm_round", args.comm_round))):
            client_indexes = self._client_sampling(g_round, int(args.client_num_in_total), int(args.client_num_per_round))
            groups = self._assign_groups(client_indexes)
            w_groups = []
            for gid, members in sorted(groups.items()):
                w_group = copy.deepcopy(w_global)
                for _ in range(int(getattr(args, "group_comm_round", 1))):
                    w_locals = []
                    for cid in members:
                        client.update_local_dataset(cid, self.train_data_local_dict[cid], self.test_data_local_dict[cid],
                                                    self.train_data_local_num_dict[cid])
                        w_locals.append((client.get_sample_number(), client.train(copy.deepcopy(w_group))))
                    w_group = A.cast_like(A.fedavg_aggregate(w_locals), w_global)
                w_groups.append((sum(self.train_data_local_num_dict[c] for c in members), w_group))
            w_global = A.cast_like(A.fedavg_aggregate(w_groups), w_global)
            self.model_trainer.set_model_params(w_global)
            freq = int(getattr(args, "frequency_of_the_test", 1) or 1)
            if g_round % freq == 0 or g_round == int(getattr(args, "global_comm_round", args.comm_round)) - 1:
                self._local_test_on_all_clients(g_round)
        return w_global
