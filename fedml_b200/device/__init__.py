"""``fedml.device`` — device selection (reference: device/device.py:8-72, device/gpu_mapping*.py).

* simulation / single_process: ``cuda:<gpu_id>`` when ``using_gpu`` else cpu;
* simulation / MPI (multi-process): YAML map ``{host: [procs_on_gpu0, procs_on_gpu1, ...]}`` ->
  process_id -> (host, gpu) with ``sum == worker_num`` checked; the reference returns
  ``torch.device("mps")`` unconditionally when a mapping file is given —
  here the mapped CUDA device is returned;
* simulation / NCCL, distributed: ``cuda:LOCAL_RANK`` (one process per GPU);
* cross-silo: ``device_type`` gpu/mps/cpu; hierarchical uses the same YAML map keyed by
  ``proc_rank_in_silo``.
"""
import logging
import os
import socket

import torch
import yaml

__all__ = ["get_device", "mapping_processes_to_gpu_device_from_yaml_file"]


def _cuda_or_cpu(index):
    if torch.cuda.is_available():
        return torch.device(f"cuda:{int(index) % torch.cuda.device_count()}")
    return torch.device("cpu")


def mapping_processes_to_gpu_device_from_yaml_file(process_id, worker_number, gpu_util_file, gpu_util_key,
                                                   check_sum=True):
    if gpu_util_file is None:
        return _cuda_or_cpu(process_id)
    with open(gpu_util_file, "r") as fh:
        table = yaml.safe_load(fh)[gpu_util_key]
    placement = []
    for host, per_gpu in table.items():
        for gpu_idx, n_proc in enumerate(per_gpu):
            placement.extend([(host, gpu_idx)] * int(n_proc))
    if check_sum and len(placement) != worker_number:
        raise AssertionError(f"gpu mapping lists {len(placement)} processes, expected {worker_number}")
    host, gpu = placement[process_id % len(placement)]
    logging.info("process %d -> host %s gpu %d", process_id, host, gpu)
    return _cuda_or_cpu(gpu)


def get_device(args):
    tt = getattr(args, "training_type", "simulation")
    backend = str(getattr(args, "backend", "single_process"))
    using_gpu = bool(getattr(args, "using_gpu", False))
    if tt == "simulation" and backend in ("single_process", "sp"):
        return _cuda_or_cpu(getattr(args, "gpu_id", 0)) if using_gpu else torch.device("cpu")
    if backend.upper() == "NCCL" or tt == "distributed":
        return _cuda_or_cpu(int(os.environ.get("LOCAL_RANK", getattr(args, "local_rank", 0) or 0)))
    if tt == "simulation":                         # multi-process simulator
        if not using_gpu:
            return torch.device("cpu")
        return mapping_processes_to_gpu_device_from_yaml_file(
            int(getattr(args, "process_id", getattr(args, "rank", 0))), int(getattr(args, "worker_num", 1)) + 1,
            getattr(args, "gpu_mapping_file", None), getattr(args, "gpu_mapping_key", None))
    if tt == "cross_silo":
        if getattr(args, "scenario", "horizontal") == "hierarchical" and using_gpu:
            return mapping_processes_to_gpu_device_from_yaml_file(
                int(getattr(args, "proc_rank_in_silo", 0)), int(getattr(args, "n_proc_in_silo", 1)),
                getattr(args, "gpu_mapping_file", None), getattr(args, "gpu_mapping_key", None), check_sum=False)
        dt = str(getattr(args, "device_type", "gpu")).lower()
        if using_gpu and dt == "gpu":
            return _cuda_or_cpu(getattr(args, "gpu_id", 0))
        if using_gpu and dt == "mps" and getattr(torch.backends, "mps", None) and torch.backends.mps.is_available():
            return torch.device("mps")
        return torch.device("cpu")
    return _cuda_or_cpu(getattr(args, "gpu_id", 0)) if using_gpu else torch.device("cpu")


def get_gpu_util_map(silo_gpu_util_file: str) -> dict:
    """``{process index: (host, gpu index)}`` from a silo's yaml ``{host: [processes on gpu 0, processes on gpu 1, ...]}``
    (reference device/gpu_mapping_cross_silo.py:69-85): processes are numbered host by host, GPU by GPU."""
    with open(silo_gpu_util_file) as fh:
        per_host = yaml.safe_load(fh) or {}
    out, i = {}, 0
    for host, per_gpu in per_host.items():
        for gpu, n in enumerate(per_gpu):
            for _ in range(int(n)):
                out[i] = (host, gpu)
                i += 1
    logging.info("GPU MAPPING: %s", out)
    return out
