"""Model files of the cross-device (Beehive) server: genuine MNN ``.mnn`` graphs, read and written without the MNN runtime.

The reference's server reads every uploaded ``.mnn`` as ``{layer_idx: tensor}`` (cross_device/server_mnn/utils.py:11-28),
averages, writes the values back into the MNN graph (:31-50) and evaluates.  :mod:`.mnn_format` does the same on the
FlatBuffers bytes (the MNN wheel cannot be installed here): parameters are read in graph order, written back in place, and
the graphs of the reference's mobile models (LeNet-5, logistic regression, the ResNet-20 training graph) are built from scratch;
other architectures are stored as a parameters-only MNN graph.  Files written by round 1 of this repo (magic ``FB2M``) are still readable."""
import os
from collections import OrderedDict

import numpy as np
import torch

from ..core.distributed.communication.message import Message
from . import mnn_format as MF

_LENET_KEYS = ["conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"]


def write_mnn_like(path, state_dict, input_dims=None):
    """``state_dict`` -> ``.mnn``: the full graph for LeNet-5, for pure linear stacks and for the CIFAR ResNet-20 (running
    statistics are not part of MNN's training graph and are left out), parameters-only otherwise."""
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    keys = list(state_dict.keys())
    arrs = [v.detach().cpu().float().numpy() for v in state_dict.values()]
    if keys == _LENET_KEYS:
        return MF.build_lenet5(path, arrs, input_dims or (1, 1, 28, 28))
    learned = [a for k, a in zip(keys, arrs) if not k.endswith(("running_mean", "running_var", "num_batches_tracked"))]
    if len(learned) == 65 and learned[0].ndim == 4 and learned[-2].ndim == 2:
        want = MF.resnet20_param_shapes(learned[-2].shape[0], learned[0].shape[1])
        if [tuple(a.shape) for a in learned] == want:
            return MF.build_resnet20(path, learned, input_dims or (1, learned[0].shape[1], 32, 32))
    linear = len(keys) % 2 == 0 and all(keys[2 * i].endswith("weight") and keys[2 * i + 1].endswith("bias") and arrs[2 * i].ndim == 2
                                        and arrs[2 * i + 1].ndim == 1 for i in range(len(keys) // 2))
    if linear and keys:
        return MF.build_mlp(path, arrs, input_dims or (1, arrs[0].shape[1]))
    return MF.build_params_only(path, OrderedDict(zip(keys, arrs)))


def _is_legacy(path):
    with open(path, "rb") as fh:
        return fh.read(4) == b"FB2M"


def read_mnn_like(path) -> "OrderedDict[str, torch.Tensor]":
    """Ordered ``name -> tensor``; names are the MNN op names (``TrainableParam<n>`` for graphs a device SDK wrote)."""
    if _is_legacy(path):
        with open(path, "rb") as fh:
            m = Message.from_bytes(fh.read())
        return OrderedDict(zip(m.get("names"), m.get("tensors")))
    return OrderedDict((k, torch.from_numpy(v)) for k, v in MF.read_params(path).items())


def read_mnn_as_tensor_dict(path):
    return {i: t for i, t in enumerate(read_mnn_like(path).values())}


def write_tensor_dict_to_mnn(path, tensor_params_dict):
    """Values of ``{layer_idx: tensor}`` into the graph at ``path``, in place (the graph itself is untouched)."""
    n = len(tensor_params_dict)
    vals = [np.asarray(tensor_params_dict[i].detach().cpu().float().numpy() if torch.is_tensor(tensor_params_dict[i])
                       else tensor_params_dict[i]) for i in range(n)]
    if _is_legacy(path):
        names = list(read_mnn_like(path).keys())
        m = Message()
        m.add_params("names", names)
        m.add_params("tensors", [torch.as_tensor(v) for v in vals])
        with open(path + ".tmp", "wb") as fh:
            fh.write(m.to_bytes())
        os.replace(path + ".tmp", path)
        return path
    return MF.write_params(path, vals)


def load_into_module(path, module):
    """Parameters of the file into ``module``: by name when the names are state-dict keys, else positionally over the
    module's parameters-and-buffers order restricted to tensors of matching element count (MNN stores linear biases as
    [1, out] and names parameters ``TrainableParam<n>``)."""
    sd = read_mnn_like(path)
    own = module.state_dict()
    if set(sd.keys()) <= set(own.keys()):
        module.load_state_dict({k: sd[k].reshape(own[k].shape).to(own[k].dtype) for k in sd}, strict=False)
        return module
    targets = [k for k, v in own.items() if v.dtype.is_floating_point and not k.endswith("num_batches_tracked")]
    vals = list(sd.values())
    if len(vals) != len(targets):                              # MNN training graphs carry no BatchNorm running statistics
        targets = [k for k, _ in module.named_parameters()]
    if len(vals) != len(targets) or any(vals[i].numel() != own[targets[i]].numel() for i in range(len(vals))):
        raise ValueError(f"{path}: {len(vals)} parameters do not line up with the model's {len(targets)} tensors")
    module.load_state_dict({k: vals[i].reshape(own[k].shape).to(own[k].dtype) for i, k in enumerate(targets)}, strict=False)
    return module
