"""YAML + CLI configuration.

Behavioural contract (reference: python/fedml/arguments.py:32-147):
  * three CLI flags: ``--yaml_config_file/--cf``, ``--run_id`` (str, "0"), ``--rank`` (int, 0);
  * every key of every YAML section is flattened onto the args object, section names
    are cosmetic;
  * when ``--cf`` is empty a packaged default YAML is chosen from (training_type, backend);
  * hierarchical cross-silo merges one more per-rank YAML (server_config_path for rank 0,
    client_silo_config_paths[rank-1] otherwise).

What is new here: a typed schema with defaults (``SCHEMA``) so that code never needs
``hasattr`` probes, ``Arguments.from_dict`` for programmatic construction (tests, bench),
and a ``blackwell_args`` family for the B200 runtime knobs.
"""
from __future__ import annotations

import argparse
import copy
import json
import os
from os import path
from typing import Any, Dict, Iterable, Optional

import yaml

from .constants import (
    FEDML_CROSS_SILO_SCENARIO_HIERARCHICAL,
    FEDML_SIMULATION_TYPE_MPI,
    FEDML_SIMULATION_TYPE_NCCL,
    FEDML_SIMULATION_TYPE_SP,
    FEDML_TRAINING_PLATFORM_CROSS_DEVICE,
    FEDML_TRAINING_PLATFORM_CROSS_SILO,
    FEDML_TRAINING_PLATFORM_DISTRIBUTED,
    FEDML_TRAINING_PLATFORM_SIMULATION,
)

_HERE = path.abspath(path.dirname(__file__))

# key -> (type, default).  ``None`` type means "leave as parsed by YAML".
SCHEMA: Dict[str, tuple] = {
    # common_args
    "training_type": (str, FEDML_TRAINING_PLATFORM_SIMULATION),
    "using_mlops": (bool, False),
    "random_seed": (int, 0),
    "scenario": (str, "horizontal"),
    "config_version": (str, "release"),
    # data_args
    "dataset": (str, "mnist"),
    "data_cache_dir": (str, "./data"),
    "partition_method": (str, "hetero"),
    "partition_alpha": (float, 0.5),
    "train_size": (int, -1),
    "test_size": (int, -1),
    "synthetic_data": (bool, True),
    # model_args
    "model": (str, "lr"),
    "model_file_cache_folder": (str, "./model_file_cache"),
    "global_model_file_path": (str, "./model_file_cache/global_model.pt"),
    # train_args
    "federated_optimizer": (str, "FedAvg"),
    "client_id_list": (None, "[]"),
    "client_num_in_total": (int, 1000),
    "client_num_per_round": (int, 10),
    "comm_round": (int, 10),
    "epochs": (int, 1),
    "batch_size": (int, 10),
    "client_optimizer": (str, "sgd"),
    "learning_rate": (float, 0.03),
    "weight_decay": (float, 0.001),
    "train_batch_size": (int, 0),
    "test_batch_size": (int, 0),
    "server_optimizer": (str, "sgd"),
    "server_lr": (float, 1.0),
    "server_momentum": (float, 0.0),
    "server_beta1": (float, 0.9),
    "server_beta2": (float, 0.99),
    "server_eps": (float, 1e-3),
    "fedprox_mu": (float, 0.0),
    "mu": (float, 0.0),
    "gmf": (float, 0.0),
    "momentum": (float, 0.0),
    "dampening": (float, 0.0),
    "nesterov": (bool, False),
    "group_method": (str, "random"),
    "group_num": (int, 2),
    "global_comm_round": (int, 2),
    "group_comm_round": (int, 2),
    # validation_args
    "frequency_of_the_test": (int, 5),
    "ci": (int, 0),
    # device_args
    "using_gpu": (bool, False),
    "gpu_id": (int, 0),
    "worker_num": (int, 1),
    "gpu_mapping_file": (None, None),
    "gpu_mapping_key": (None, None),
    "server_gpu_mapping_key": (None, None),
    "device_type": (str, "gpu"),
    # comm_args
    "backend": (str, FEDML_SIMULATION_TYPE_SP),
    "is_mobile": (int, 0),
    "mqtt_config_path": (None, None),
    "s3_config_path": (None, None),
    "trpc_master_config_path": (None, None),
    "grpc_ipconfig_path": (None, None),
    "round_deadline_s": (float, 0.0),
    # tracking_args
    "log_file_dir": (str, "./log"),
    "enable_wandb": (bool, False),
    "wandb_key": (None, None),
    "wandb_entity": (None, None),
    "wandb_project": (str, "fedml"),
    "wandb_name": (str, "fedml_b200"),
    "run_name": (str, "fedml_b200"),
    "local_log_output_path": (None, None),
    # attack / defence
    "attack_case": (None, None),
    "defense_type": (None, None),
    "norm_bound": (float, 5.0),
    "stddev": (float, 0.025),
    "poison_type": (None, None),
    "attack_freq": (int, 10),
    "byzantine_fraction": (float, 0.0),
    "byzantine_attack": (str, "sign_flip"),
    # blackwell_args (new)
    "clients_per_gpu_batch": (int, 0),
    "agg_dtype": (str, "fp32"),
    "compute_dtype": (str, "bf16"),
    "use_multicast": (bool, True),
    "use_cuda_graphs": (bool, True),
    "fused_sgd_restage": (bool, True),      # plain SGD: optimizer step fused with re-staging the bf16 operands
    "client_groups": (int, 0),              # 0 = auto: clients of a GPU advance in this many concurrent stream groups
    "sparse_embedding_sgd": (bool, True),   # plain SGD: update only the word-embedding rows of tokens that occur
    "dp_clip": (float, 0.0),
    "dp_sigma": (float, 0.0),
    "robust_agg": (str, "none"),
    "krum_f": (int, 0),
    "krum_m": (int, 1),
    "trim_ratio": (float, 0.1),
    "checkpoint_dir": (None, None),
    "checkpoint_every": (int, 0),
    "resume": (bool, False),
}


def _coerce(key: str, val: Any) -> Any:
    spec = SCHEMA.get(key)
    if spec is None or spec[0] is None or val is None:
        return val
    typ = spec[0]
    try:
        if typ is bool:
            if isinstance(val, str):
                return val.strip().lower() in ("1", "true", "yes", "on")
            return bool(val)
        return typ(val)
    except (TypeError, ValueError):
        return val


def add_args(argv: Optional[Iterable[str]] = None) -> argparse.Namespace:
    parser = argparse.ArgumentParser(description="fedml_b200")
    parser.add_argument("--yaml_config_file", "--cf", type=str, default="",
                        help="yaml configuration file")
    parser.add_argument("--run_id", type=str, default="0")
    parser.add_argument("--rank", type=int, default=0)
    # unknown flags are tolerated: launchers (torchrun, pytest) add their own.
    args, _ = parser.parse_known_args(list(argv) if argv is not None else None)
    return args


def default_config_path(training_type, comm_backend) -> str:
    if training_type == FEDML_TRAINING_PLATFORM_SIMULATION:
        sub = {
            FEDML_SIMULATION_TYPE_SP: "simulation_sp",
            FEDML_SIMULATION_TYPE_MPI: "simulation_mpi",
            FEDML_SIMULATION_TYPE_NCCL: "simulation_nccl",
        }.get(comm_backend, "simulation_sp")
    elif training_type == FEDML_TRAINING_PLATFORM_CROSS_SILO:
        sub = "cross_silo"
    elif training_type == FEDML_TRAINING_PLATFORM_CROSS_DEVICE:
        sub = "cross_device"
    elif training_type == FEDML_TRAINING_PLATFORM_DISTRIBUTED:
        sub = "distributed"
    else:
        sub = "simulation_sp"
    return path.join(_HERE, "config", sub, "fedml_config.yaml")


class Arguments:
    """Flat attribute bag holding CLI flags plus every YAML key."""

    def __init__(self, cmd_args=None, training_type=None, comm_backend=None,
                 apply_defaults: bool = True):
        if apply_defaults:
            for k, (_, d) in SCHEMA.items():
                setattr(self, k, copy.deepcopy(d))
        if cmd_args is None:
            cmd_args = argparse.Namespace(yaml_config_file="", run_id="0", rank=0)
        for k, v in vars(cmd_args).items():
            setattr(self, k, v)
        self.yaml_paths = []
        self._load(cmd_args, training_type, comm_backend)

    # -- construction helpers -------------------------------------------------
    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "Arguments":
        self = cls.__new__(cls)
        for k, (_, dv) in SCHEMA.items():
            setattr(self, k, copy.deepcopy(dv))
        self.yaml_config_file, self.run_id, self.rank, self.yaml_paths = "", "0", 0, []
        for k, v in d.items():
            setattr(self, k, _coerce(k, v))
        return self

    @staticmethod
    def load_yaml_config(yaml_path: str) -> Dict[str, Any]:
        with open(yaml_path, "r") as stream:
            try:
                return yaml.safe_load(stream) or {}
            except yaml.YAMLError as exc:
                raise ValueError(f"Yaml error - check yaml file {yaml_path}: {exc}")

    def set_attr_from_config(self, configuration: Dict[str, Any]) -> None:
        for family, params in configuration.items():
            if isinstance(params, dict):
                for key, val in params.items():
                    setattr(self, key, _coerce(key, val))
            else:  # tolerate flat YAMLs
                setattr(self, family, _coerce(family, params))

    def _load(self, cmd_args, training_type, comm_backend) -> None:
        cfg = getattr(cmd_args, "yaml_config_file", "") or ""
        used_default = cfg == ""
        if used_default:
            cfg = default_config_path(training_type, comm_backend)
            self.yaml_config_file = cfg
        self.yaml_paths = [cfg]
        configuration = self.load_yaml_config(cfg)
        self.set_attr_from_config(configuration)
        if used_default and training_type == FEDML_TRAINING_PLATFORM_SIMULATION \
                and comm_backend == FEDML_SIMULATION_TYPE_MPI:
            self.gpu_mapping_file = path.join(_HERE, "config", "simulation_mpi", "gpu_mapping.yaml")
        tt = training_type or getattr(self, "training_type", None)
        if tt == FEDML_TRAINING_PLATFORM_CROSS_SILO and \
                getattr(self, "scenario", None) == FEDML_CROSS_SILO_SCENARIO_HIERARCHICAL:
            extra = None
            if int(self.rank) == 0 and getattr(self, "server_config_path", None):
                extra = self.server_config_path
            elif getattr(self, "client_silo_config_paths", None):
                extra = self.client_silo_config_paths[int(self.rank) - 1]
            if extra:
                base = path.dirname(path.abspath(cfg))
                if not path.isabs(extra) and not path.exists(extra):
                    extra = path.join(base, extra)
                self.yaml_paths.append(extra)
                self.set_attr_from_config(self.load_yaml_config(extra))

    # -- conveniences ----------------------------------------------------------
    def get(self, key: str, default=None):
        return getattr(self, key, default)

    def client_ids(self):
        """``client_id_list`` arrives as a JSON string in YAML ("[1, 2]")."""
        v = getattr(self, "client_id_list", "[]")
        if isinstance(v, str):
            try:
                v = json.loads(v) if v.strip() else []
            except json.JSONDecodeError:
                v = []
        return list(v or [])

    def to_dict(self) -> Dict[str, Any]:
        return {k: v for k, v in vars(self).items() if not k.startswith("_")}

    def __repr__(self) -> str:  # short, for logs
        keys = ("training_type", "backend", "federated_optimizer", "dataset", "model",
                "client_num_in_total", "client_num_per_round", "comm_round")
        return "Arguments(" + ", ".join(f"{k}={getattr(self, k, None)!r}" for k in keys) + ")"


def load_arguments(training_type=None, comm_backend=None, argv=None) -> Arguments:
    cmd_args = add_args(argv)
    return Arguments(cmd_args, training_type, comm_backend)


def substitute_fedsys_vars(text: str, mapping: Dict[str, Any]) -> str:
    """``${FEDSYS.KEY}`` templating used by MLOps packages (reference: cli/edge_deployment/login.py:79-108)."""
    for k, v in mapping.items():
        text = text.replace("${FEDSYS.%s}" % k, str(v))
    return text
