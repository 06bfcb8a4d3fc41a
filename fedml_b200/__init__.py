"""fedml_b200 — a Blackwell-native federated-learning framework with FedML's API surface.

Public entry points mirror the reference package (python/fedml/__init__.py:34-304):
``init``, ``run_simulation``, ``run_cross_silo_server/client``,
``run_hierarchical_cross_silo_server/client``, ``run_mnn_server``, ``run_distributed`` (a stub in the
reference; the Cheetah data-parallel trainer here), ``load_arguments``, the ``ClientTrainer`` /
``ServerAggregator`` ABCs and the ``device`` / ``data`` / ``model`` / ``simulation`` / ``cross_silo`` /
``cross_device`` sub-packages, plus ``FedMLRunner`` (upstream FedML's entry object).

Nothing heavy is imported here: the reference hard-requires mpi4py, MNN, paho, boto3, wandb,
wget and h5py at ``import fedml`` time; every optional dependency in this
package is imported lazily at its point of use.
"""
import logging
import os
import random

import numpy as np
import torch

from . import constants
from .arguments import Arguments, load_arguments
from .constants import (FEDML_SIMULATION_TYPE_MPI, FEDML_SIMULATION_TYPE_NCCL, FEDML_SIMULATION_TYPE_SP,
                        FEDML_TRAINING_PLATFORM_CROSS_DEVICE, FEDML_TRAINING_PLATFORM_CROSS_SILO,
                        FEDML_TRAINING_PLATFORM_DISTRIBUTED, FEDML_TRAINING_PLATFORM_SIMULATION)
from .core.alg_frame import ClientTrainer, ServerAggregator
from .runner import FedMLRunner

__version__ = "0.7.39+b200.1"

_LAZY = {"device", "data", "model", "models", "simulation", "cross_silo", "cross_device", "core", "ops", "parallel",
         "utils", "cli", "distributed"}


def __getattr__(name):
    import importlib
    if name == "model":
        mod = importlib.import_module(".models", __name__)
    elif name == "distributed":
        mod = importlib.import_module(".parallel.cheetah", __name__)
    elif name in _LAZY:
        mod = importlib.import_module("." + name, __name__)
    else:
        raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
    globals()[name] = mod
    return mod


def seed_everything(seed: int):
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    torch.backends.cudnn.deterministic = True


def init(args=None):
    """Load the YAML config (``--cf``), set up logging, seed every RNG, fill in rank plumbing."""
    if args is None:
        args = load_arguments()
    from .utils.logging import init_logs
    init_logs(args)
    seed_everything(int(getattr(args, "random_seed", 0)))
    if getattr(args, "enable_wandb", False):
        try:
            import wandb
            os.environ.setdefault("WANDB_MODE", "offline")
            wandb.init(project=getattr(args, "wandb_project", "fedml"), name=getattr(args, "wandb_name", None),
                       config=args.to_dict() if hasattr(args, "to_dict") else None)
        except Exception as exc:  # offline box
            logging.warning("wandb unavailable: %s", exc)
    tt = getattr(args, "training_type", FEDML_TRAINING_PLATFORM_SIMULATION)
    backend = getattr(args, "backend", FEDML_SIMULATION_TYPE_SP)
    env = os.environ
    if tt == FEDML_TRAINING_PLATFORM_SIMULATION and backend == FEDML_SIMULATION_TYPE_MPI:
        # the reference reads MPI.COMM_WORLD here; our multi-process simulator gets the same three
        # attributes from the launcher environment (torchrun-style) or the --rank flag
        args.process_id = int(env.get("RANK", getattr(args, "rank", 0)))
        args.worker_num = int(env.get("WORLD_SIZE", int(getattr(args, "worker_num", 1)) + 1)) - 1 \
            if "WORLD_SIZE" in env else int(getattr(args, "worker_num", 1))
        args.comm = None
    elif tt == FEDML_TRAINING_PLATFORM_SIMULATION and backend == FEDML_SIMULATION_TYPE_NCCL:
        args.process_id = int(env.get("RANK", 0))
        args.local_rank = int(env.get("LOCAL_RANK", 0))
        args.world_size = int(env.get("WORLD_SIZE", 1))
    elif tt == FEDML_TRAINING_PLATFORM_CROSS_SILO:
        if getattr(args, "scenario", "horizontal") == "hierarchical":
            # torchrun plumbing inside a silo (reference __init__.py:78-130)
            args.n_proc_in_silo = int(env.get("WORLD_SIZE", getattr(args, "n_proc_in_silo", 1) or 1))
            args.proc_rank_in_silo = int(env.get("RANK", getattr(args, "proc_rank_in_silo", 0) or 0))
            args.rank_in_node = int(env.get("LOCAL_RANK", 0))
            args.pg_master_address = env.get("MASTER_ADDR", getattr(args, "pg_master_address", "127.0.0.1"))
            args.pg_master_port = int(env.get("MASTER_PORT", getattr(args, "pg_master_port", 29500) or 29500))
        else:
            args.n_proc_in_silo, args.proc_rank_in_silo = 1, 0
        args.process_id = int(getattr(args, "rank", 0))
        ids = args.client_ids() if hasattr(args, "client_ids") else []
        if int(getattr(args, "rank", 0)) > 0 and ids and int(args.rank) <= len(ids):
            args.edge_id = ids[int(args.rank) - 1]
        elif not hasattr(args, "edge_id"):
            args.edge_id = int(getattr(args, "rank", 0))
    elif tt == FEDML_TRAINING_PLATFORM_CROSS_DEVICE:
        args.rank = 0
        args.process_id = 0
    elif tt == FEDML_TRAINING_PLATFORM_DISTRIBUTED:
        args.process_id = int(env.get("RANK", 0))
        args.local_rank = int(env.get("LOCAL_RANK", 0))
        args.world_size = int(env.get("WORLD_SIZE", 1))
    return args


def _bootstrap(training_type, backend=None, argv=None):
    from . import data as _data, device as _device, models as _models
    args = init(load_arguments(training_type, backend, argv))
    if backend is not None and training_type == FEDML_TRAINING_PLATFORM_SIMULATION:
        args.backend = backend
    device = _device.get_device(args)
    loader = _data.load_cross_silo if training_type == FEDML_TRAINING_PLATFORM_CROSS_SILO else _data.load
    dataset, output_dim = loader(args)
    model = _models.create(args, output_dim)
    return args, device, dataset, model


def run_simulation(backend=FEDML_SIMULATION_TYPE_SP, argv=None):
    """One-line API: init -> device -> data -> model -> simulator.run() (reference :139-169)."""
    from .simulation import SimulatorMPI, SimulatorNCCL, SimulatorSingleProcess
    args, device, dataset, model = _bootstrap(FEDML_TRAINING_PLATFORM_SIMULATION, backend, argv)
    cls = {FEDML_SIMULATION_TYPE_SP: SimulatorSingleProcess, FEDML_SIMULATION_TYPE_MPI: SimulatorMPI,
           FEDML_SIMULATION_TYPE_NCCL: SimulatorNCCL}.get(backend)
    if cls is None:
        raise ValueError(f"unknown simulation backend {backend!r}")
    return cls(args, device, dataset, model).run()


def run_cross_silo_server(argv=None):
    from .cross_silo import Server
    args, device, dataset, model = _bootstrap(FEDML_TRAINING_PLATFORM_CROSS_SILO, None, argv)
    return Server(args, device, dataset, model).run()


def run_cross_silo_client(argv=None):
    from .cross_silo import Client
    args, device, dataset, model = _bootstrap(FEDML_TRAINING_PLATFORM_CROSS_SILO, None, argv)
    return Client(args, device, dataset, model).run()


def run_hierarchical_cross_silo_server(argv=None):
    from .cross_silo.hierarchical import Server
    args, device, dataset, model = _bootstrap(FEDML_TRAINING_PLATFORM_CROSS_SILO, None, argv)
    return Server(args, device, dataset, model).run()


def run_hierarchical_cross_silo_client(argv=None):
    from .cross_silo.hierarchical import Client
    args, device, dataset, model = _bootstrap(FEDML_TRAINING_PLATFORM_CROSS_SILO, None, argv)
    return Client(args, device, dataset, model).run()


def run_mnn_server(argv=None):
    from .cross_device import ServerMNN
    args, device, dataset, model = _bootstrap(FEDML_TRAINING_PLATFORM_CROSS_DEVICE, None, argv)
    return ServerMNN(args, device, dataset[3], model).run()


def run_distributed(argv=None):
    """Cheetah: data-parallel training with the fused gradient all-reduce (reference: ``pass``)."""
    from .parallel.cheetah import CheetahRunner
    args, device, dataset, model = _bootstrap(FEDML_TRAINING_PLATFORM_DISTRIBUTED, None, argv)
    return CheetahRunner(args, device, dataset, model).run()


__all__ = ["init", "run_simulation", "run_cross_silo_server", "run_cross_silo_client",
           "run_hierarchical_cross_silo_server", "run_hierarchical_cross_silo_client", "run_mnn_server",
           "run_distributed", "load_arguments", "Arguments", "ClientTrainer", "ServerAggregator", "FedMLRunner",
           "constants", "seed_everything", "__version__"]
