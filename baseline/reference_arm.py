"""Reference arm of bench.py: the UNMODIFIED reference (pip-installed into baseline/_ref) driven through
its own public API — ``fedml.simulation.SimulatorSingleProcess(args, device, dataset, model).run()`` with the
reference's own ResNet-56 (``fedml.model.create``), trainer, client loop and CPU aggregation.

Nothing from fedml_b200 is imported here.  What this harness adds around the reference:
  * import stubs for third-party packages the reference imports at module load but that are not
    installable offline (mpi4py, MNN, paho, boto3, wget, h5py, wandb internals) — none of them is used
    on the single-process FedAvg path;
  * synthetic CIFAR-10-shaped client data (the reference loader downloads CIFAR-10, impossible offline),
    built as per-client ``DataLoader(shuffle=True, drop_last=True)`` exactly like
    ``data/cifar10/data_loader.py:209-214`` and passed as the fork's 9-item dataset list;
  * evaluation hooks are neutralised on the instance (``_local_test_on_all_clients`` and
    ``_validate_global_model``) because the metric counts training rounds only, for both arms;
  * round boundaries are time-stamped by wrapping ``_client_sampling`` (called once at the start of a round).
The reference has no working multi-GPU simulator in this snapshot (SimulatorNCCL is a stub, the MPI
simulator needs mpi4py), so for N > 1 only rank 0 runs and the result is flagged ``single_gpu_only``.
"""
import importlib.abc
import importlib.machinery
import json
import os
import sys
import time
import types
from unittest import mock

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")

_STUB_ROOTS = ("mpi4py", "MNN", "paho", "boto3", "botocore", "wget", "h5py")


class _StubLoader(importlib.abc.Loader):
    def create_module(self, spec):
        m = types.ModuleType(spec.name)
        m.__path__ = []
        m.__getattr__ = lambda name: mock.MagicMock(name=f"{spec.name}.{name}")
        return m

    def exec_module(self, module):
        pass


class _StubFinder(importlib.abc.MetaPathFinder):
    def __init__(self, extra=()):
        self.extra = set(extra)

    def find_spec(self, fullname, path=None, target=None):
        root = fullname.split(".")[0]
        if root in _STUB_ROOTS or fullname in self.extra:
            return importlib.machinery.ModuleSpec(fullname, _StubLoader(), is_package=True)
        return None


def import_reference():
    if HERE not in sys.path:
        sys.path.insert(0, HERE)
    from install_reference import ensure_reference
    ensure_reference()          # installs from /root/reference/python when baseline/_ref is not there yet
    # the reference ships protoc<3.19 generated gRPC stubs; the pure-python protobuf runtime still accepts
    # them (must be set before google.protobuf is first imported)
    os.environ.setdefault("PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION", "python")
    sys.meta_path.append(_StubFinder())
    # `wandb.sdk.internal.stats` was removed from newer wandb; the reference imports it for MLOps system stats
    try:
        import wandb.sdk.internal.stats  # noqa: F401
    except Exception:
        sys.meta_path.append(_StubFinder(extra={"wandb.sdk.internal.stats"}))
    os.environ.setdefault("WANDB_MODE", "disabled")
    sys.path.insert(0, REF)
    for k in [k for k in sys.modules if k == "fedml" or k.startswith("fedml.")]:
        del sys.modules[k]
    import fedml  # the reference package
    assert os.path.abspath(fedml.__file__).startswith(REF), fedml.__file__
    return fedml


def make_dataset_flat(torch, np, n_clients, per_client, batch_size, dim=784, seed=0):
    """MNIST-shaped clients (flattened 28x28 floats in [0, 1]) for the logistic-regression config."""
    g = np.random.default_rng(seed)
    from torch.utils.data import DataLoader, TensorDataset
    train_local, test_local, nums = {}, {}, {}
    xt = torch.from_numpy(g.random((batch_size, dim), dtype=np.float32))
    yt = torch.from_numpy(g.integers(0, 10, batch_size).astype(np.int64))
    test_loader = DataLoader(TensorDataset(xt, yt), batch_size=batch_size)
    for c in range(n_clients):
        x = torch.from_numpy(g.random((per_client, dim), dtype=np.float32))
        y = torch.from_numpy(g.integers(0, 10, per_client).astype(np.int64))
        train_local[c] = DataLoader(TensorDataset(x, y), batch_size=batch_size, shuffle=True, drop_last=False)
        test_local[c] = test_loader
        nums[c] = per_client
    return [n_clients * per_client, batch_size, test_loader, test_loader, nums, train_local, test_local, 10, None]


def make_dataset(torch, np, n_clients, per_client, batch_size, seed=0):
    g = np.random.default_rng(seed)
    mean = np.array([0.4914, 0.4822, 0.4465], dtype=np.float32).reshape(1, 3, 1, 1)
    std = np.array([0.2470, 0.2435, 0.2616], dtype=np.float32).reshape(1, 3, 1, 1)
    from torch.utils.data import DataLoader, TensorDataset
    train_local, test_local, nums = {}, {}, {}
    xt = torch.from_numpy(((g.integers(0, 256, (batch_size, 3, 32, 32)).astype(np.float32) / 255 - mean) / std))
    yt = torch.from_numpy(g.integers(0, 10, batch_size).astype(np.int64))
    test_loader = DataLoader(TensorDataset(xt, yt), batch_size=batch_size)
    for c in range(n_clients):
        x = torch.from_numpy((g.integers(0, 256, (per_client, 3, 32, 32)).astype(np.float32) / 255 - mean) / std)
        y = torch.from_numpy(g.integers(0, 10, per_client).astype(np.int64))
        train_local[c] = DataLoader(TensorDataset(x, y), batch_size=batch_size, shuffle=True, drop_last=True)
        test_local[c] = test_loader
        nums[c] = per_client
    n_train = n_clients * per_client
    return [n_train, batch_size, test_loader, test_loader, nums, train_local, test_local, 10, None]


def run(cfg, steps, warmup, n_gpus):
    import numpy as np
    import torch
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None            # the reference SP simulator drives a single GPU
    fedml = import_reference()
    import logging
    logging.disable(logging.INFO)
    args = types.SimpleNamespace(
        training_type="simulation", backend="single_process", dataset=cfg.get("dataset", "cifar10"), model=cfg["model"],
        federated_optimizer="FedAvg", client_num_in_total=cfg["client_num_in_total"],
        client_num_per_round=cfg["client_num_per_round"], comm_round=warmup + steps, epochs=cfg["epochs"],
        batch_size=cfg["batch_size"], client_optimizer="sgd", learning_rate=cfg["learning_rate"], weight_decay=0.001,
        frequency_of_the_test=10 ** 9, using_gpu=True, gpu_id=0, random_seed=0, enable_wandb=False, rank=0, run_id="0",
        using_mlops=False)
    torch.manual_seed(0); np.random.seed(0)
    on_cpu = bool(cfg.get("cpu"))
    device = torch.device("cuda:0" if (torch.cuda.is_available() and not on_cpu) else "cpu")
    if on_cpu:
        args.using_gpu = False
        dataset = make_dataset_flat(torch, np, cfg["client_num_in_total"], cfg["samples_per_client"], cfg["batch_size"])
    else:
        dataset = make_dataset(torch, np, cfg["client_num_in_total"], cfg["samples_per_client"], cfg["batch_size"])
    model = fedml.model.create(args, 10)
    from fedml.simulation import SimulatorSingleProcess
    sim = SimulatorSingleProcess(args, device, dataset, model)
    api = sim.fl_trainer
    api._local_test_on_all_clients = lambda *a, **k: {}
    api._validate_global_model = lambda *a, **k: (0.0, 0.0)
    stamps = []
    orig = api._client_sampling

    def stamped(*a, **k):
        if device.type == "cuda":
            torch.cuda.synchronize()
        stamps.append(time.perf_counter())
        return orig(*a, **k)

    api._client_sampling = stamped
    import joblib
    with mock.patch.object(joblib, "dump", lambda *a, **k: None):
        sim.run()
    if device.type == "cuda":
        torch.cuda.synchronize()
    stamps.append(time.perf_counter())
    per_round = [b - a for a, b in zip(stamps[:-1], stamps[1:])][warmup:]
    ms = 1e3 * sum(per_round) / len(per_round)
    K = cfg["client_num_per_round"]
    return {"ms_per_step": ms, "rounds_per_sec": 1e3 / ms, "client_updates_per_sec": K * 1e3 / ms}
