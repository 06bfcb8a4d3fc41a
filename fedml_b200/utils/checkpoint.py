"""Checkpoint import/export and resumable run state.

Format compatibility: a model checkpoint is a ``torch.save``-pickled
``OrderedDict`` whose keys are the module paths of the model (incl. int64
``num_batches_tracked``), optionally wrapped as ``{state_dict, optim_dict, epoch, ...}``
(FedGKT ``last.pth``/``best.pth``, GKTServerTrainer.py:262-290); loaders accept a ``module.``
prefix (resnet.py:265-277).  The reference has no resume path at all; ``save_run_state`` /
``load_run_state`` add one: round index, global model, server-optimizer moments, NumPy / torch
RNG state and DP RNG counters, written atomically.
"""
import os
import shutil
from collections import OrderedDict

import numpy as np
import torch


def strip_module_prefix(sd):
    return OrderedDict((k[7:] if k.startswith("module.") else k, v) for k, v in sd.items())


def save_state_dict(path, state_dict):
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    cpu = OrderedDict((k, v.detach().cpu()) for k, v in state_dict.items())
    tmp = path + ".tmp"
    torch.save(cpu, tmp)
    os.replace(tmp, path)


def load_state_dict(path):
    obj = torch.load(path, map_location="cpu", weights_only=False)
    if isinstance(obj, dict) and "state_dict" in obj and not torch.is_tensor(obj["state_dict"]):
        obj = obj["state_dict"]
    return strip_module_prefix(obj)


def save_checkpoint(state, is_best, checkpoint_dir, filename="last.pth"):
    """FedGKT-style ``{state_dict, optim_dict, epoch, test_accTop1, test_accTop5}`` + ``best.pth`` copy."""
    os.makedirs(checkpoint_dir, exist_ok=True)
    path = os.path.join(checkpoint_dir, filename)
    torch.save(state, path)
    if is_best:
        shutil.copyfile(path, os.path.join(checkpoint_dir, "best.pth"))
    return path


def save_run_state(ckpt_dir, next_round, state_dict, extra=None):
    os.makedirs(ckpt_dir, exist_ok=True)
    blob = {"round": int(next_round),
            "state_dict": OrderedDict((k, v.detach().cpu()) for k, v in state_dict.items()),
            "extra": extra or {},
            "rng": {"numpy": np.random.get_state(), "torch": torch.get_rng_state(),
                    "cuda": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else None}}
    tmp = os.path.join(ckpt_dir, "run_state.pt.tmp")
    torch.save(blob, tmp)
    os.replace(tmp, os.path.join(ckpt_dir, "run_state.pt"))
    save_state_dict(os.path.join(ckpt_dir, "global_model.pt"), state_dict)


def load_run_state(ckpt_dir, restore_rng=True):
    path = os.path.join(ckpt_dir, "run_state.pt")
    if not os.path.exists(path):
        return None
    blob = torch.load(path, map_location="cpu", weights_only=False)
    if restore_rng and "rng" in blob:
        np.random.set_state(blob["rng"]["numpy"])
        torch.set_rng_state(blob["rng"]["torch"])
        if blob["rng"].get("cuda") is not None and torch.cuda.is_available():
            try:
                torch.cuda.set_rng_state_all(blob["rng"]["cuda"])
            except Exception:
                pass
    return blob
