"""Default client trainers (eager PyTorch path; also the training oracle for the batched kernels).

Numerics follow the reference (``my_model_trainer_classification.py:18-91``,
``my_model_trainer_nwp.py``, ``my_model_trainer_tag_prediction.py``): a fresh optimizer per call;
``sgd`` is ``SGD(lr)`` with no momentum and no weight decay (weight decay only reaches Adam, which
uses ``amsgrad=True``); ``epochs`` passes over the pre-built batches.  Additions: a real FedProx
proximal term ``mu/2 * ||w - w_global||^2`` (the reference's fedprox directory has none, SURVEY
§0.2), the FedNova local optimizer, and parameters that stay on the training device (callers
that need CPU copies ask for them explicitly instead of paying a D2H per client per round).
"""
from __future__ import annotations

import copy
import logging
from collections import OrderedDict

import numpy as np
import torch
import torch.nn as nn

from ..core.alg_frame import ClientTrainer


def _to_device_batch(x, y, device):
    return x.to(device, non_blocking=True), y.to(device, non_blocking=True)


def build_client_optimizer(params, args):
    name = str(getattr(args, "client_optimizer", "sgd")).lower()
    lr = float(getattr(args, "learning_rate", getattr(args, "lr", 0.03)))
    if name == "sgd":
        return torch.optim.SGD(params, lr=lr)
    if name == "sgd_momentum":
        return torch.optim.SGD(params, lr=lr, momentum=float(getattr(args, "momentum", 0.9)),
                               weight_decay=float(getattr(args, "weight_decay", 0.0)))
    return torch.optim.Adam(params, lr=lr, weight_decay=float(getattr(args, "weight_decay", 0.0)), amsgrad=True)


class ModelTrainerCLS(ClientTrainer):
    """Classification trainer: CrossEntropy, optional FedProx term."""

    criterion_factory = staticmethod(lambda: nn.CrossEntropyLoss())
    keep_on_device = True

    def get_model_params(self):
        if self.keep_on_device:
            return OrderedDict((k, v.detach().clone()) for k, v in self.model.state_dict().items())
        return self.model.cpu().state_dict()

    def set_model_params(self, model_parameters):
        self.model.load_state_dict(model_parameters)

    # -- hooks for subclasses ------------------------------------------------
    def _loss(self, criterion, logits, labels):
        return criterion(logits, labels)

    def train(self, train_data, device, args):
        model = self.model.to(device)
        model.train()
        cw = getattr(args, "class_weight", None)           # per-client class weights (S-FedAvg: s_fedavg/fedavg_api.py:112-136)
        criterion = (nn.CrossEntropyLoss(weight=torch.as_tensor(cw, dtype=torch.float32).to(device)) if cw is not None
                     else self.criterion_factory()).to(device)
        trainable = [p for p in model.parameters() if p.requires_grad]
        optimizer = build_client_optimizer(trainable, args)
        mu = float(getattr(args, "fedprox_mu", 0.0) or 0.0)
        if str(getattr(args, "federated_optimizer", "")).lower() == "fedprox" and mu == 0.0:
            mu = float(getattr(args, "mu", 0.0) or 0.0)
        anchor = [p.detach().clone() for p in trainable] if mu > 0 else None
        clip = float(getattr(args, "client_grad_clip", 0.0) or 0.0)
        epoch_loss = []
        for _ in range(int(getattr(args, "epochs", 1))):
            losses = []
            for x, labels in train_data:
                x, labels = _to_device_batch(x, labels, device)
                optimizer.zero_grad(set_to_none=True)
                loss = self._loss(criterion, model(x), labels)
                if anchor is not None:
                    loss = loss + 0.5 * mu * sum(((p - a) ** 2).sum() for p, a in zip(trainable, anchor))
                loss.backward()
                if clip > 0:
                    torch.nn.utils.clip_grad_norm_(trainable, clip)
                optimizer.step()
                losses.append(loss.detach())
            if losses:
                epoch_loss.append(float(torch.stack(losses).mean()))   # one sync per epoch, not per step
        self.last_train_loss = epoch_loss[-1] if epoch_loss else float("nan")
        return self.last_train_loss

    def test(self, test_data, device, args):
        model = self.model.to(device)
        model.eval()
        criterion = self.criterion_factory().to(device)
        correct = torch.zeros((), device=device)
        loss_sum = torch.zeros((), device=device)
        total = 0
        ncls = None
        tp = fp = fn = None
        with torch.no_grad():
            for x, target in test_data:
                x, target = _to_device_batch(x, target, device)
                logits = model(x)
                loss_sum += self._loss(criterion, logits, target) * target.size(0)
                pred = logits.argmax(dim=1)
                correct += (pred == target).sum()
                total += target.size(0)
                if ncls is None:
                    ncls = logits.shape[1]
                    tp = torch.zeros(ncls, device=device); fp = torch.zeros_like(tp); fn = torch.zeros_like(tp)
                hit = pred == target
                tp += torch.bincount(target[hit], minlength=ncls)[:ncls]
                fp += torch.bincount(pred[~hit], minlength=ncls)[:ncls]
                fn += torch.bincount(target[~hit], minlength=ncls)[:ncls]
        metrics = {"test_correct": float(correct), "test_loss": float(loss_sum), "test_total": total}
        if ncls is not None:
            # per-class recall / precision dicts: the fork's additions (fedavg/my_model_trainer_classification.py:101-153)
            rec = (tp / (tp + fn).clamp_min(1)).tolist(); pre = (tp / (tp + fp).clamp_min(1)).tolist()
            metrics["test_recall"] = {i: r for i, r in enumerate(rec)}
            metrics["test_precision"] = {i: p for i, p in enumerate(pre)}
        return metrics

    def test_on_the_server(self, train_data_local_dict, test_data_local_dict, device, args=None) -> bool:
        return False


class ModelTrainerNWP(ModelTrainerCLS):
    """Next-word prediction: CE with ``ignore_index=0`` over [B, V, S] logits."""
    criterion_factory = staticmethod(lambda: nn.CrossEntropyLoss(ignore_index=0))

    def test(self, test_data, device, args):
        model = self.model.to(device); model.eval()
        crit = self.criterion_factory().to(device)
        correct = loss_sum = 0.0
        total = 0
        with torch.no_grad():
            for x, target in test_data:
                x, target = _to_device_batch(x, target, device)
                logits = model(x)
                loss_sum += float(crit(logits, target)) * target.size(0)
                pred = logits.argmax(dim=1)
                mask = target != 0 if target.dim() > 1 else torch.ones_like(target, dtype=torch.bool)
                correct += float(((pred == target) & mask).sum())
                total += int(mask.sum())
        return {"test_correct": correct, "test_loss": loss_sum, "test_total": total}


class ModelTrainerTAGPred(ModelTrainerCLS):
    """Multi-label tag prediction: ``BCELoss(reduction='sum')`` on sigmoid outputs."""
    criterion_factory = staticmethod(lambda: nn.BCELoss(reduction="sum"))

    def test(self, test_data, device, args):
        model = self.model.to(device); model.eval()
        crit = self.criterion_factory().to(device)
        m = {"test_correct": 0.0, "test_loss": 0.0, "test_precision": 0.0, "test_recall": 0.0, "test_total": 0}
        with torch.no_grad():
            for x, target in test_data:
                x, target = _to_device_batch(x, target, device)
                pred = model(x)
                m["test_loss"] += float(crit(pred, target))
                hard = (pred > 0.5).float()
                m["test_correct"] += float((hard == target).all(dim=1).sum())
                tp = (hard * target).sum(dim=1)
                m["test_precision"] += float((tp / hard.sum(dim=1).clamp_min(1e-13)).sum())
                m["test_recall"] += float((tp / target.sum(dim=1).clamp_min(1e-13)).sum())
                m["test_total"] += target.size(0)
        return m


def create_model_trainer(model, args):
    ds = str(getattr(args, "dataset", "")).lower()
    if ds == "stackoverflow_lr":
        return ModelTrainerTAGPred(model, args)
    if ds in ("fed_shakespeare", "stackoverflow_nwp"):
        return ModelTrainerNWP(model, args)
    return ModelTrainerCLS(model, args)


# --------------------------------------------------------------------------- FedNova local optimizer
class FedNovaOptimizer(torch.optim.Optimizer):
    """SGD + momentum + proximal term that also tracks the normalising scalar ``a_i`` and the
    cumulative update (reference: single_process/fednova/fednova.py:96-169, with the removed
    ``Tensor.add_(scalar, tensor)`` overloads restated).

    per step: ``d = g (+wd*w) -> momentum buffer -> (+mu*(w - w0))``; ``cum_grad += lr*d``; ``w -= lr*d``.
    normaliser: momentum != 0: ``c <- c*m + 1; a += c``;  ``lr*mu != 0``: ``a <- a*(1 - lr*mu) + 1``;
    otherwise ``a += 1``.
    """

    def __init__(self, params, lr, ratio=1.0, gmf=0.0, mu=0.0, momentum=0.0, dampening=0.0, weight_decay=0.0,
                 nesterov=False):
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                        nesterov=nesterov)
        super().__init__(params, defaults)
        self.ratio, self.gmf, self.mu, self.momentum = ratio, gmf, mu, momentum
        self.local_normalizing_vec = 0.0
        self.local_counter = 0.0
        self.local_steps = 0
        self.etamu = 0.0

    @torch.no_grad()
    def step(self, closure=None):
        for group in self.param_groups:
            lr, wd, mom, damp = group["lr"], group["weight_decay"], group["momentum"], group["dampening"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                d = p.grad
                if wd:
                    d = d.add(p, alpha=wd)
                st = self.state[p]
                if "old_init" not in st:
                    st["old_init"] = p.detach().clone()
                if mom:
                    buf = st.get("momentum_buffer")
                    if buf is None:
                        buf = st["momentum_buffer"] = d.detach().clone()
                    else:
                        buf.mul_(mom).add_(d, alpha=1 - damp)
                    d = d.add(buf, alpha=mom) if group["nesterov"] else buf
                if self.mu:
                    d = d.add(p - st["old_init"], alpha=self.mu)
                if "cum_grad" not in st:
                    st["cum_grad"] = d.detach().clone().mul_(lr)
                else:
                    st["cum_grad"].add_(d, alpha=lr)
                p.add_(d, alpha=-lr)
            self.etamu = lr * self.mu
        if self.momentum:
            self.local_counter = self.local_counter * self.momentum + 1
            self.local_normalizing_vec += self.local_counter
        if self.etamu:
            self.local_normalizing_vec = self.local_normalizing_vec * (1 - self.etamu) + 1
        if not self.momentum and not self.etamu:
            self.local_normalizing_vec += 1
        self.local_steps += 1
