"""Re-emits every fedml_config.yaml of the package and of examples/ in the documented house style.

    python tools/annotate_configs.py            # rewrite in place
    python tools/annotate_configs.py --check    # exit 1 if a file is not in the house style (used by the tests)

House style: sections in the order a run consumes them (what kind of run -> where it runs -> how workers talk -> Blackwell switches
-> data -> model -> optimisation -> evaluation -> logging -> back-end specific sections), one key per line with a trailing comment
saying what the key does *in this framework*.  The values are kept exactly; the schema (section and key names) is the one
``fedml_b200.arguments.load_arguments`` flattens, which is also the reference's (python/fedml/arguments.py:60-118).
"""
from __future__ import annotations

import glob
import os
import sys

import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SECTION_ORDER = ["common_args", "device_args", "comm_args", "blackwell_args", "data_args", "model_args", "train_args",
                 "validation_args", "tracking_args"]

SECTION_DOC = {
    "common_args": "what kind of run this is",
    "device_args": "where it runs",
    "comm_args": "how server and workers exchange models",
    "blackwell_args": "B200 execution switches (all optional; defaults in fedml_b200/arguments.py)",
    "data_args": "dataset and how it is cut into clients",
    "model_args": "network",
    "train_args": "federated optimisation",
    "validation_args": "evaluation cadence",
    "tracking_args": "logs and experiment tracking",
    "mqtt_args": "MQTT broker + payload store (python -m fedml_b200.core.distributed.communication.broker starts the bundled broker)",
    "grpc_args": "gRPC back-end: rank -> ip table and first port",
    "trpc_args": "torch.distributed.rpc back-end",
    "hierarchical_args": "cross-silo hierarchical: every silo is a torchrun job of n_node_in_silo x n_proc_per_node processes",
}

KEY_DOC = {
    "training_type": "simulation | cross_silo | cross_device | distributed  (selects the runner)",
    "scenario": "horizontal (one process per silo) | hierarchical (a DDP group per silo)",
    "using_mlops": "true: report status / metrics to an MLOps endpoint",
    "random_seed": "seeds python, numpy, torch and the client sampler",
    "using_gpu": "false runs the PyTorch fp32 path on the host",
    "gpu_id": "cuda device of this process (single-process runs)",
    "worker_num": "number of worker processes next to the server",
    "gpu_mapping_file": "yaml of host -> processes per GPU",
    "gpu_mapping_key": "entry of gpu_mapping_file to use",
    "backend": "single_process | MPI | NCCL | MQTT_S3 | MQTT_S3_MNN | GRPC | TRPC",
    "is_mobile": "1: payloads are .mnn files for the phone SDK instead of torch state dicts",
    "compute_dtype": "tf32 (reference accuracy) | bf16 | fp32 for the client-batched kernels",
    "agg_dtype": "accumulator precision of the weighted average",
    "use_multicast": "reduce through NVSwitch multimem when the fabric offers it",
    "use_cuda_graphs": "capture the local-SGD step once and replay it",
    "clients_per_gpu_batch": "clients trained side by side in one launch; 0 = as many as fit",
    "dataset": "name understood by fedml_b200.data.load",
    "data_cache_dir": "real files are read from here; absent -> synthetic data of the same shape",
    "partition_method": "hetero = Dirichlet label skew, homo = IID",
    "partition_alpha": "Dirichlet concentration (smaller = more skew)",
    "model": "name understood by fedml_b200.models.create",
    "model_file_cache_folder": "scratch directory for exchanged model files",
    "global_model_file_path": "where the server keeps the current global model",
    "federated_optimizer": "FedAvg | FedOpt | FedProx | FedNova | ... (see docs/COMPONENTS.md)",
    "client_id_list": "ids this process plays; \"[]\" = take them from the rank",
    "client_num_in_total": "clients the dataset is partitioned into",
    "client_num_per_round": "clients sampled every round",
    "comm_round": "number of communication rounds",
    "epochs": "local epochs per round",
    "batch_size": "local mini-batch",
    "client_optimizer": "sgd | adam on the clients",
    "learning_rate": "client learning rate",
    "weight_decay": "client L2 coefficient",
    "momentum": "client SGD momentum",
    "defense_type": "norm_diff_clipping | weak_dp | none",
    "norm_bound": "clip updates whose L2 distance to the global model exceeds this",
    "stddev": "gaussian noise of weak_dp",
    "attack_freq": "a poisoned client joins every attack_freq rounds",
    "poison_type": "which backdoor set the attacker trains on",
    "frequency_of_the_test": "evaluate every this many rounds (and on the last one)",
    "log_file_dir": "per-process log files go here",
    "enable_wandb": "true needs the wandb package and a key",
    "mqtt_host": "broker address",
    "mqtt_port": "1883 plain, 8883 with mqtt_tls: true",
    "object_store_dir": "shared directory standing in for the S3 bucket",
    "mqtt_binary": "true: binary tensor codec on the MQTT topics instead of JSON lists (not readable by the reference's SDK)",
    "grpc_ipconfig_path": "csv of receiver_id,ip",
    "grpc_base_port": "rank r listens on grpc_base_port + r",
    "trpc_master_config_path": "csv holding master ip and port",
    "enable_cuda_rpc": "send cuda tensors device to device",
    "n_node_in_silo": "nodes of one silo",
    "n_proc_per_node": "processes (GPUs) per node",
    "master_address": "rendezvous address inside the silo",
    "launcher_rdzv_port": "rendezvous port inside the silo",
}


def _scalar(v) -> str:
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, str):
        plain = v and all(ch.isalnum() or ch in "_-./" for ch in v) and v.lower() not in ("true", "false", "null", "yes", "no")
        return v if plain else '"' + v.replace('"', '\\"') + '"'
    if v is None:
        return "null"
    return repr(v) if isinstance(v, float) else str(v)


def render(cfg: dict, title: str) -> str:
    sections = [s for s in SECTION_ORDER if s in cfg] + [s for s in cfg if s not in SECTION_ORDER]
    out = [f"# {title}", "# layout: tools/annotate_configs.py; every key may also be given on the command line as --key value", ""]
    for s in sections:
        body = cfg[s]
        doc = SECTION_DOC.get(s)
        out.append(f"{s}:" + (f"    # {doc}" if doc else ""))
        if not isinstance(body, dict):
            out[-1] = f"{s}: {_scalar(body)}"
            continue
        width = max(len(f"  {k}: {_scalar(v)}") for k, v in body.items() if not isinstance(v, (dict, list))) if any(
            not isinstance(v, (dict, list)) for v in body.values()) else 0
        for k, v in body.items():
            if isinstance(v, (dict, list)):
                dumped = yaml.safe_dump({k: v}, default_flow_style=False, sort_keys=False).rstrip().splitlines()
                out.extend("  " + ln for ln in dumped)
                continue
            line = f"  {k}: {_scalar(v)}"
            doc = KEY_DOC.get(k)
            out.append(line.ljust(width + 3) + f"# {doc}" if doc else line)
        out.append("")
    return "\n".join(out).rstrip() + "\n"


def _title(path: str) -> str:
    rel = os.path.relpath(path, ROOT)
    d = os.path.dirname(rel)
    if rel.startswith("examples"):
        return f"{d}: run with  python {d}/main.py --cf {rel}" if os.path.exists(os.path.join(ROOT, d, "main.py")) else d
    return f"packaged default of the {os.path.basename(d)} runner: used when no --cf is given (fedml_b200/arguments.py)"


def main(check: bool) -> int:
    files = sorted(glob.glob(os.path.join(ROOT, "examples", "**", "fedml_config.yaml"), recursive=True) +
                   glob.glob(os.path.join(ROOT, "fedml_b200", "config", "**", "fedml_config.yaml"), recursive=True))
    stale = 0
    for p in files:
        with open(p) as fh:
            text = fh.read()
        cfg = yaml.safe_load(text)
        new = render(cfg, _title(p))
        assert yaml.safe_load(new) == cfg, p
        if new != text:
            stale += 1
            if check:
                print("not in house style:", os.path.relpath(p, ROOT))
            else:
                with open(p, "w") as fh:
                    fh.write(new)
                print("rewrote", os.path.relpath(p, ROOT))
    return 1 if (check and stale) else 0


if __name__ == "__main__":
    sys.exit(main("--check" in sys.argv))
