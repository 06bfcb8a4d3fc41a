"""Small helpers the reference keeps in per-algorithm ``utils.py`` files, gathered in one place.

* flat parameter views and running averages of the FedGKT trainers (mpi_p2p_mp/fedgkt/utils.py:8-160),
* weight-decay-free BatchNorm parameter groups (same file, ``bnwd_optim_params``),
* bookkeeping of the FedSeg runs (mpi_p2p_mp/fedseg/utils.py: metric record, pickle / json dumps, ``str_to_bool`` for argparse),
* the flatten -> collective -> unflatten pattern of the FedNova optimiser (single_process/fednova/comm_helpers.py:7-63),
* the logistic helper and accuracy counter of the vertical-FL fixture (single_process/classical_vertical_fl).
"""
from __future__ import annotations

import json
import pickle
from typing import Iterable, List, Sequence, Tuple

import numpy as np
import torch


# ---- flat parameters ---------------------------------------------------------------------------------------------------------
def get_flat_params_from(model: torch.nn.Module) -> torch.Tensor:
    return torch.cat([p.data.reshape(-1) for p in model.parameters()])


def set_flat_params_to(model: torch.nn.Module, flat_params: torch.Tensor) -> None:
    total = sum(p.numel() for p in model.parameters())
    if total != flat_params.numel():
        raise ValueError(f"flat vector has {flat_params.numel()} entries, the model {total}")
    offset = 0
    for p in model.parameters():
        n = p.numel()
        p.data.copy_(flat_params[offset:offset + n].view_as(p))
        offset += n


def flatten_tensors(tensors: Sequence[torch.Tensor]) -> torch.Tensor:
    """One contiguous 1-D copy of dense tensors of one dtype."""
    return tensors[0].reshape(-1).clone() if len(tensors) == 1 else torch.cat([t.reshape(-1) for t in tensors])


def unflatten_tensors(flat: torch.Tensor, tensors: Iterable[torch.Tensor]) -> Tuple[torch.Tensor, ...]:
    """Views into ``flat`` with the shapes of ``tensors``."""
    out, offset = [], 0
    for t in tensors:
        out.append(flat.narrow(0, offset, t.numel()).view_as(t))
        offset += t.numel()
    return tuple(out)


def communicate(tensors: List[torch.Tensor], communication_op) -> None:
    """Runs ``communication_op(tensor=<flat buffer>)`` (e.g. a partial of ``dist.all_reduce``) once for the whole list and copies
    the result back into the tensors."""
    flat = flatten_tensors(tensors)
    communication_op(tensor=flat)
    with torch.no_grad():
        for piece, t in zip(unflatten_tensors(flat, tensors), tensors):
            t.copy_(piece)


# ---- optimisation ------------------------------------------------------------------------------------------------------------
def split_bn_params(model: torch.nn.Module, model_params, master_params):
    """``(batch-norm master params, the other master params)``: ``model_params[i]`` says which group ``master_params[i]`` is in."""
    bn_ids = {id(p) for m in model.modules() if isinstance(m, torch.nn.modules.batchnorm._BatchNorm) for p in m.parameters()}
    pairs = list(zip(model_params, master_params))
    return [mast for mod, mast in pairs if id(mod) in bn_ids], [mast for mod, mast in pairs if id(mod) not in bn_ids]


def bnwd_optim_params(model, model_params, master_params):
    """Parameter groups that exempt BatchNorm scales / shifts from weight decay (arXiv:1807.11205)."""
    bn, rest = split_bn_params(model, model_params, master_params)
    return [{"params": bn, "weight_decay": 0}, {"params": rest}]


class RunningAverage:
    def __init__(self):
        self.steps, self.total = 0, 0

    def update(self, val):
        self.total += val
        self.steps += 1

    def value(self):
        return self.total / float(self.steps)

    __call__ = value


# ---- files -------------------------------------------------------------------------------------------------------------------
def get_state_dict(file):
    """``torch.load`` onto the CPU regardless of where the checkpoint was saved."""
    return torch.load(file, map_location="cpu", weights_only=False)


def save_dict_to_json(d: dict, json_path: str) -> None:
    with open(json_path, "w") as fh:
        json.dump({k: (float(v) if isinstance(v, (np.floating, np.integer)) else v) for k, v in d.items()}, fh, indent=4)


def save_as_pickle_file(path: str, data) -> None:
    with open(path, "wb") as fh:
        pickle.dump(data, fh, protocol=pickle.HIGHEST_PROTOCOL)


def load_from_pickle_file(path: str):
    with open(path, "rb") as fh:
        return pickle.load(fh)


def str_to_bool(s) -> bool:
    """For ``argparse`` ``type=``: "true" / "false" in any case."""
    if isinstance(s, bool):
        return s
    v = str(s).strip().lower()
    if v in ("true", "t", "1", "yes", "y"):
        return True
    if v in ("false", "f", "0", "no", "n"):
        return False
    raise ValueError(f"not a valid boolean string: {s!r}")


def count_parameters(model: torch.nn.Module) -> int:
    return sum(p.numel() for p in model.parameters() if p.requires_grad)


class EvaluationMetricsKeeper:
    """Record of one segmentation evaluation (pixel accuracy, per-class accuracy, mean IoU, frequency-weighted IoU, loss)."""

    def __init__(self, accuracy, accuracy_class, mIoU, FWIoU, loss):  # noqa: N803
        self.acc, self.acc_class, self.mIoU, self.FWIoU, self.loss = accuracy, accuracy_class, mIoU, FWIoU, loss


# ---- vertical FL -------------------------------------------------------------------------------------------------------------
def sigmoid(x):
    return 1.0 / (1.0 + np.exp(-np.asarray(x, dtype=np.float64)))


def compute_correct_prediction(*, y_targets, y_prob_preds, threshold=0.5):
    """``(thresholded predictions, [positive predictions, negative predictions, correct predictions])``."""
    y_hat = (np.asarray(y_prob_preds).reshape(-1) > threshold).astype(int)
    y = np.asarray(y_targets).reshape(-1)
    pos = int(y_hat.sum())
    return y_hat, [pos, int(len(y_hat) - pos), int((y_hat == y).sum())]
