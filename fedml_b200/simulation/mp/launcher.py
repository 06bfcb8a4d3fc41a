"""Runs a worker-per-process algorithm either as real processes (one per rank, launched by torchrun /
``--rank``; transport TCP, the mpirun replacement) or — on a single machine without a launcher — as threads in
one process over the in-proc transport.  Dispatch on ``args.federated_optimizer`` mirrors ``SimulatorMPI``
(simulation/simulator.py:43-97) and wires the algorithms the reference leaves as ``pass``."""
import copy
import os
import threading

from ...constants import canonical_optimizer_name


def _entry(name):
    from . import fedavg as F
    if name == "FedAvg_robust":
        from .fedavg_robust import FedML_FedAvgRobust_distributed
        return FedML_FedAvgRobust_distributed
    if name in ("FedAvg", "FedOpt", "FedProx"):
        return F.FedML_FedAvg_distributed
    if name == "base_framework":
        from .templates import FedML_Base_distributed
        return FedML_Base_distributed
    if name == "decentralized_fl":
        from .templates import FedML_Decentralized_Demo_distributed
        return FedML_Decentralized_Demo_distributed
    if name == "split_nn":
        from .split_nn import SplitNN_distributed
        return SplitNN_distributed
    if name == "classical_vertical":
        from .vertical_fl import FedML_VFL_distributed
        return FedML_VFL_distributed
    if name == "FedGKT":
        from .fedgkt import FedML_FedGKT_distributed
        return FedML_FedGKT_distributed
    if name == "FedNAS":
        from .fednas import FedML_FedNAS_distributed
        return FedML_FedNAS_distributed
    if name == "FedGAN":
        from .fedgan import FedML_FedGan_distributed
        return FedML_FedGan_distributed
    if name == "FedSeg":
        from .fedseg import FedML_FedSeg_distributed
        return FedML_FedSeg_distributed
    if name == "turbo_aggregate":
        from .turboaggregate import FedML_TurboAggregate_distributed
        return FedML_TurboAggregate_distributed
    raise ValueError(f"federated_optimizer {name!r} is not available in the multi-process simulator")


class _ThreadedSimulation:
    def __init__(self, entry, args, device, dataset, model, model_trainer, server_aggregator):
        self.entry, self.args, self.device, self.dataset, self.model = entry, args, device, dataset, model
        self.model_trainer, self.server_aggregator = model_trainer, server_aggregator
        self.results = {}

    def run(self):
        W = int(getattr(self.args, "worker_num", 0) or self.args.client_num_per_round)
        size = W + 1
        from ...core.distributed.communication.inproc import InProcHub
        InProcHub.reset(getattr(self.args, "run_id", 0))
        errors = []

        def work(rank):
            try:
                a = copy.copy(self.args)
                a.rank = a.process_id = rank
                m = copy.deepcopy(self.model)
                kw = {}
                if rank == 0 and self.server_aggregator is not None:
                    kw["server_aggregator"] = self.server_aggregator
                self.results[rank] = self.entry(a, rank, size, None, self.device, self.dataset, m,
                                                copy.deepcopy(self.model_trainer) if self.model_trainer else None,
                                                backend="INPROC", **kw)
            except BaseException as exc:  # noqa: BLE001
                errors.append((rank, exc))
                raise

        # every worker thread that runs CPU kernels gets its own intra-op (OpenMP) team: size the teams so that all ranks together
        # use the cores once instead of `size` times (restored afterwards; `cpu_threads_per_rank` overrides)
        import torch
        saved_threads = torch.get_num_threads()
        on_cpu = str(getattr(self.device, "type", self.device)) == "cpu"
        per_rank = int(getattr(self.args, "cpu_threads_per_rank", 0) or 0) or (max(1, saved_threads // size) if on_cpu else saved_threads)
        torch.set_num_threads(per_rank)
        try:
            return self._run_threads(work, size, errors)
        finally:
            torch.set_num_threads(saved_threads)

    def _run_threads(self, work, size, errors):
        from ...core.distributed.communication.inproc import InProcHub
        threads = [threading.Thread(target=work, args=(r,), daemon=True) for r in range(0, size)]
        for t in threads:
            t.start()
        import time
        deadline = time.time() + float(getattr(self.args, "simulation_timeout_s", 3600))
        # fail fast: one crashed rank would leave the others blocked in their receive loops forever
        while any(t.is_alive() for t in threads):
            if errors:
                InProcHub.reset(getattr(self.args, "run_id", 0))
                raise RuntimeError(f"rank {errors[0][0]} failed") from errors[0][1]
            if time.time() > deadline:
                raise TimeoutError("multi-process simulation (threaded) timed out")
            time.sleep(0.02)
        if errors:
            raise RuntimeError(f"rank {errors[0][0]} failed") from errors[0][1]
        return self.results.get(0)


class _ProcessSimulation:
    def __init__(self, entry, args, device, dataset, model, model_trainer, server_aggregator):
        self.fn = lambda: entry(args, int(args.process_id), int(args.worker_num) + 1, None, device, dataset, model,
                                model_trainer, backend=getattr(args, "mp_transport", "TCP"),
                                **({"server_aggregator": server_aggregator} if int(args.process_id) == 0 and server_aggregator else {}))

    def run(self):
        return self.fn()


def build_mp_simulation(args, device, dataset, model, model_trainer=None, server_aggregator=None):
    entry = _entry(canonical_optimizer_name(args.federated_optimizer))
    multi_process = "WORLD_SIZE" in os.environ and int(os.environ["WORLD_SIZE"]) > 1
    if multi_process:
        args.process_id = int(os.environ.get("RANK", getattr(args, "rank", 0)))
        args.worker_num = int(os.environ["WORLD_SIZE"]) - 1
        return _ProcessSimulation(entry, args, device, dataset, model, model_trainer, server_aggregator)
    return _ThreadedSimulation(entry, args, device, dataset, model, model_trainer, server_aggregator)
