"""FedSeg — segmentation FedAvg (reference: mpi_p2p_mp/fedseg/*.py, utils.py:66-298): clients also report
evaluation metrics (pixel acc, class acc, mIoU, FWIoU, loss); ``SegmentationLosses`` (CE / focal),
``LR_Scheduler`` (poly / cos / step), ``Evaluator`` confusion matrix, ``Saver`` checkpoints."""
import math
import os

import numpy as np
import torch
import torch.nn as nn

from ...core.alg_frame import ClientTrainer
from .fedavg import FedML_FedAvg_distributed


class SegmentationLosses:
    def __init__(self, size_average=True, batch_average=True, ignore_index=255):
        self.ignore_index, self.size_average, self.batch_average = ignore_index, size_average, batch_average

    def build_loss(self, mode="ce"):
        return self.CrossEntropyLoss if mode == "ce" else self.FocalLoss

    def CrossEntropyLoss(self, logit, target):  # noqa: N802
        loss = nn.functional.cross_entropy(logit, target.long(), ignore_index=self.ignore_index,
                                           reduction="mean" if self.size_average else "sum")
        return loss / logit.shape[0] if (self.batch_average and not self.size_average) else loss

    def FocalLoss(self, logit, target, gamma=2, alpha=0.5):  # noqa: N802
        logpt = -nn.functional.cross_entropy(logit, target.long(), ignore_index=self.ignore_index, reduction="mean")
        pt = torch.exp(logpt)
        return -((1 - pt) ** gamma) * alpha * logpt


class LR_Scheduler:  # noqa: N801
    def __init__(self, mode, base_lr, num_epochs, iters_per_epoch=1, lr_step=0):
        self.mode, self.lr, self.N, self.ipe, self.lr_step = mode, base_lr, num_epochs * iters_per_epoch, iters_per_epoch, lr_step

    def __call__(self, optimizer, i, epoch):
        T = epoch * self.ipe + i
        if self.mode == "cos":
            lr = 0.5 * self.lr * (1 + math.cos(1.0 * T / self.N * math.pi))
        elif self.mode == "poly":
            lr = self.lr * pow((1 - 1.0 * T / self.N), 0.9)
        else:
            lr = self.lr * (0.1 ** (epoch // max(self.lr_step, 1)))
        for g in optimizer.param_groups:
            g["lr"] = lr
        return lr


class Evaluator:
    def __init__(self, num_class):
        self.num_class = num_class
        self.confusion_matrix = np.zeros((num_class,) * 2)

    def add_batch(self, gt, pred):
        mask = (gt >= 0) & (gt < self.num_class)
        label = self.num_class * gt[mask].astype(int) + pred[mask]
        self.confusion_matrix += np.bincount(label, minlength=self.num_class ** 2).reshape(self.num_class, self.num_class)

    def Pixel_Accuracy(self):  # noqa: N802
        return np.diag(self.confusion_matrix).sum() / max(self.confusion_matrix.sum(), 1)

    def Pixel_Accuracy_Class(self):  # noqa: N802
        with np.errstate(divide="ignore", invalid="ignore"):
            return np.nanmean(np.diag(self.confusion_matrix) / self.confusion_matrix.sum(axis=1))

    def Mean_Intersection_over_Union(self):  # noqa: N802
        cm = self.confusion_matrix
        with np.errstate(divide="ignore", invalid="ignore"):
            return np.nanmean(np.diag(cm) / (cm.sum(1) + cm.sum(0) - np.diag(cm)))

    def Frequency_Weighted_Intersection_over_Union(self):  # noqa: N802
        cm = self.confusion_matrix
        with np.errstate(divide="ignore", invalid="ignore"):
            freq = cm.sum(1) / cm.sum()
            iu = np.diag(cm) / (cm.sum(1) + cm.sum(0) - np.diag(cm))
        return (freq[freq > 0] * iu[freq > 0]).sum()

    def reset(self):
        self.confusion_matrix[:] = 0


class Saver:
    def __init__(self, args):
        base = os.path.join("run", str(getattr(args, "dataset", "seg")), str(getattr(args, "model", "segnet")),
                            str(getattr(args, "checkname", "fedseg")))
        runs = sorted(d for d in (os.listdir(base) if os.path.isdir(base) else []) if d.startswith("experiment_"))
        self.experiment_dir = os.path.join(base, f"experiment_{len(runs)}")
        os.makedirs(self.experiment_dir, exist_ok=True)

    def save_checkpoint(self, state, is_best, filename="checkpoint.pth.tar"):
        path = os.path.join(self.experiment_dir, filename)
        torch.save(state, path)
        if is_best:
            with open(os.path.join(self.experiment_dir, "best_pred.txt"), "w") as fh:
                fh.write(str(state.get("best_pred", 0.0)))
            torch.save(state, os.path.join(self.experiment_dir, "model_best.pth.tar"))
        return path


class SegTrainer(ClientTrainer):
    def get_model_params(self):
        return {k: v.detach().clone() for k, v in self.model.state_dict().items()}

    def set_model_params(self, p):
        self.model.load_state_dict(p)

    def train(self, train_data, device, args):
        m = self.model.to(device); m.train()
        crit = SegmentationLosses().build_loss(str(getattr(args, "loss_type", "ce")))
        opt = torch.optim.SGD(m.parameters(), lr=float(getattr(args, "learning_rate", 0.05)), momentum=0.9)
        sched = LR_Scheduler(str(getattr(args, "lr_scheduler", "poly")), float(getattr(args, "learning_rate", 0.05)),
                             int(getattr(args, "epochs", 1)), max(len(train_data), 1))
        for ep in range(int(getattr(args, "epochs", 1))):
            for i, (x, y) in enumerate(train_data):
                sched(opt, i, ep)
                opt.zero_grad(set_to_none=True)
                crit(m(x.to(device)), y.to(device)).backward()
                opt.step()

    def test(self, test_data, device, args):
        m = self.model.to(device); m.eval()
        ev, loss, n = None, 0.0, 0
        crit = SegmentationLosses().build_loss("ce")
        with torch.no_grad():
            for x, y in test_data:
                out = m(x.to(device))
                ev = ev or Evaluator(out.shape[1])
                loss += float(crit(out, y.to(device))) * x.shape[0]; n += x.shape[0]
                ev.add_batch(y.numpy(), out.argmax(1).cpu().numpy())
        acc = ev.Pixel_Accuracy() if ev else 0.0
        return {"test_correct": acc * n, "test_loss": loss, "test_total": n, "test_acc_class": ev.Pixel_Accuracy_Class() if ev else 0,
                "test_mIoU": ev.Mean_Intersection_over_Union() if ev else 0, "test_FWIoU": ev.Frequency_Weighted_Intersection_over_Union() if ev else 0}


def FedML_FedSeg_distributed(args, process_id, worker_number, comm, device, dataset, model, model_trainer=None,  # noqa: N802
                             backend="INPROC", **kw):
    return FedML_FedAvg_distributed(args, process_id, worker_number, comm, device, dataset, model,
                                    model_trainer or SegTrainer(model, args), backend=backend, **kw)
