"""Portable on-disk model container used where the reference uses ``.mnn`` files.

The cross-device server of the reference reads every uploaded ``.mnn`` as ``{layer_idx: tensor}``
(cross_device/server_mnn/utils.py:11-28), averages, writes the values back into the MNN graph (:31-50) and
evaluates with the MNN runtime.  MNN 1.1.6 is a third-party binary wheel that cannot be installed here, so
files named ``*.mnn`` in this framework hold the same information — an ordered list of named tensors — in the
``Message`` binary codec (magic ``FB2M``).  ``read_mnn_as_tensor_dict`` keeps the reference's
index-keyed view; a real MNN file is rejected with a clear error."""
import os
from collections import OrderedDict

import torch

from ..core.distributed.communication.message import Message


def write_mnn_like(path, state_dict):
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    m = Message()
    m.add_params("names", list(state_dict.keys()))
    m.add_params("tensors", [v.detach().cpu() for v in state_dict.values()])
    tmp = path + ".tmp"
    with open(tmp, "wb") as fh:
        fh.write(m.to_bytes())
    os.replace(tmp, path)
    return path


def read_mnn_like(path) -> "OrderedDict[str, torch.Tensor]":
    with open(path, "rb") as fh:
        blob = fh.read()
    if blob[:4] != b"FB2M":
        raise ValueError(f"{path} is not a fedml_b200 model container (a genuine MNN graph needs the MNN runtime, "
                         "which is unavailable on this box)")
    m = Message.from_bytes(blob)
    return OrderedDict(zip(m.get("names"), m.get("tensors")))


def read_mnn_as_tensor_dict(path):
    return {i: t for i, t in enumerate(read_mnn_like(path).values())}


def write_tensor_dict_to_mnn(path, tensor_params_dict):
    sd = read_mnn_like(path)
    names = list(sd.keys())
    return write_mnn_like(path, OrderedDict((names[i], tensor_params_dict[i]) for i in range(len(names))))
