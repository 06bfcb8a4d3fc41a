"""Scenario runners of the Parrot simulator (reference: simulation/simulator.py:28-108).

``SimulatorSingleProcess`` dispatches on ``args.federated_optimizer`` — including the algorithms
the reference ships but never wires (FedOpt, FedNova, hierarchical, decentralized, vertical,
TurboAggregate).  ``SimulatorMPI`` keeps the name but runs the worker-per-process
protocol over this package's own transports (no mpi4py).  ``SimulatorNCCL`` — a stub in the
reference (``self.simulator = None``) — is the multi-GPU product path: clients sharded over
GPUs, batched local training, and the fused sm_100a round kernel.
"""
from ..constants import canonical_optimizer_name as _canon


def _sp_api(name):
    from .sp import fedavg_api as F
    table = {"FedAvg": F.FedAvgAPI, "FedOpt": F.FedOptAPI, "FedProx": F.FedProxAPI, "FedNova": F.FedNovaAPI,
             "FedAvg_robust": F.FedAvgRobustAPI, "HierarchicalFL": F.HierarchicalFLAPI}
    if name in table:
        return table[name]
    if name == "decentralized_fl":
        from .sp.decentralized import DecentralizedFLAPI
        return DecentralizedFLAPI
    if name == "classical_vertical":
        from .sp.vertical_fl import VerticalFLAPI
        return VerticalFLAPI
    if name == "turbo_aggregate":
        from .sp.turboaggregate import TurboAggregateAPI
        return TurboAggregateAPI
    if name in ("S_FedAvg", "HS_FedAvg"):
        from .sp.shapley_fedavg import HSFedAvgAPI, SFedAvgAPI
        return SFedAvgAPI if name == "S_FedAvg" else HSFedAvgAPI
    raise ValueError(f"federated_optimizer {name!r} is not available in the single-process simulator")


class SimulatorSingleProcess:
    def __init__(self, args, device, dataset, model, model_trainer=None, server_aggregator=None):
        api = _sp_api(_canon(args.federated_optimizer))
        try:
            self.fl_trainer = api(args, device, dataset, model, model_trainer, server_aggregator)
        except TypeError:
            self.fl_trainer = api(args, device, dataset, model)

    def run(self):
        return self.fl_trainer.train()


class SimulatorMPI:
    def __init__(self, args, device, dataset, model, model_trainer=None, server_aggregator=None):
        from .mp.launcher import build_mp_simulation
        self.simulator = build_mp_simulation(args, device, dataset, model, model_trainer, server_aggregator)

    def run(self):
        return self.simulator.run()


class SimulatorNCCL:
    def __init__(self, args, device, dataset, model, model_trainer=None, server_aggregator=None):
        from .nccl.simulator import NCCLSimulator
        self.simulator = NCCLSimulator(args, device, dataset, model, model_trainer, server_aggregator)

    def run(self):
        return self.simulator.train()
