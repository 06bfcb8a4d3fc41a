"""Centralized baseline trainer (reference: centralized/centralized_trainer.py:9-164 and
examples/centralized/main.py:495-609): a plain epoch loop over ``train_data_global``; with
``data_parallel == 1`` the model is wrapped for multi-process data parallelism (``DistributedSampler.set_epoch``
in the reference, :50-51) — here either DDP or, on GPUs, the Cheetah engine (fused gradient all-reduce)."""
import logging

import torch
import torch.nn as nn

from ..utils.metrics import MetricsSink


class CentralizedTrainer:
    def __init__(self, dataset, model, device, args):
        self.device, self.args = device, args
        (self.train_num, self.test_num, self.train_global, self.test_global) = dataset[:4]
        self.model = model.to(device)
        self.criterion = nn.CrossEntropyLoss().to(device)
        name = str(getattr(args, "client_optimizer", "sgd")).lower()
        lr, wd = float(getattr(args, "learning_rate", getattr(args, "lr", 0.03))), float(getattr(args, "weight_decay", 0.0))
        self.optimizer = torch.optim.SGD(self.model.parameters(), lr=lr, weight_decay=wd,
                                         momentum=float(getattr(args, "momentum", 0.0))) if name == "sgd" else \
            torch.optim.Adam(self.model.parameters(), lr=lr, weight_decay=wd, amsgrad=True)
        self.metrics = MetricsSink(args)
        self.history = []
        if int(getattr(args, "data_parallel", 0)) == 1 and torch.distributed.is_initialized():
            from torch.nn.parallel import DistributedDataParallel as DDP
            self.model = DDP(self.model, device_ids=[device.index] if device.type == "cuda" else None)

    def train(self):
        for epoch in range(int(getattr(self.args, "epochs", 1))):
            sampler = getattr(self.train_global, "sampler", None)
            if int(getattr(self.args, "data_parallel", 0)) == 1 and hasattr(sampler, "set_epoch"):
                sampler.set_epoch(epoch)
            self.train_impl(epoch)
            self.eval_impl(epoch)
        return self.history

    def train_impl(self, epoch_idx):
        self.model.train()
        for x, y in self.train_global:
            x, y = x.to(self.device), y.to(self.device)
            self.optimizer.zero_grad(set_to_none=True)
            loss = self.criterion(self.model(x), y)
            loss.backward()
            self.optimizer.step()

    def eval_impl(self, epoch_idx):
        freq = int(getattr(self.args, "frequency_of_train_acc_report", 1) or 1)
        stats = {"epoch": epoch_idx}
        for name, data in (("train", self.train_global), ("test", self.test_global)):
            if name == "train" and epoch_idx % freq:
                continue
            c = l = n = 0.0
            self.model.eval()
            with torch.no_grad():
                for x, y in data:
                    x, y = x.to(self.device), y.to(self.device)
                    o = self.model(x)
                    l += float(self.criterion(o, y)) * y.size(0); c += float((o.argmax(1) == y).sum()); n += y.size(0)
            stats[f"{name}_acc"], stats[f"{name}_loss"] = c / max(n, 1), l / max(n, 1)
        self.history.append(stats)
        self.metrics.log({"Train/Acc": stats.get("train_acc"), "Train/Loss": stats.get("train_loss"),
                          "Test/Acc": stats.get("test_acc"), "Test/Loss": stats.get("test_loss"), "epoch": epoch_idx})
        return stats
