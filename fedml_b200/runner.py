"""``FedMLRunner`` — the upstream-FedML entry object (absent in the 0.7.39 fork).

``FedMLRunner(args, device, dataset, model, client_trainer=None, server_aggregator=None).run()``
dispatches on ``args.training_type`` / ``args.backend`` / ``args.scenario`` to the same runners the
legacy one-liners (``fedml.run_simulation`` etc.) use.
"""
from .constants import (FEDML_CROSS_SILO_SCENARIO_HIERARCHICAL, FEDML_SIMULATION_TYPE_MPI, FEDML_SIMULATION_TYPE_NCCL,
                        FEDML_SIMULATION_TYPE_SP, FEDML_TRAINING_PLATFORM_CROSS_DEVICE,
                        FEDML_TRAINING_PLATFORM_CROSS_SILO, FEDML_TRAINING_PLATFORM_DISTRIBUTED,
                        FEDML_TRAINING_PLATFORM_SIMULATION)


class FedMLRunner:
    def __init__(self, args, device, dataset, model, client_trainer=None, server_aggregator=None):
        self.args = args
        tt = getattr(args, "training_type", FEDML_TRAINING_PLATFORM_SIMULATION)
        if tt == FEDML_TRAINING_PLATFORM_SIMULATION:
            self.runner = self._init_simulation_runner(args, device, dataset, model, client_trainer, server_aggregator)
        elif tt == FEDML_TRAINING_PLATFORM_CROSS_SILO:
            self.runner = self._init_cross_silo_runner(args, device, dataset, model, client_trainer, server_aggregator)
        elif tt == FEDML_TRAINING_PLATFORM_CROSS_DEVICE:
            from .cross_device import ServerMNN
            self.runner = ServerMNN(args, device, dataset, model, server_aggregator)
        elif tt == FEDML_TRAINING_PLATFORM_DISTRIBUTED:
            from .parallel.cheetah import CheetahRunner
            self.runner = CheetahRunner(args, device, dataset, model)
        else:
            raise ValueError(f"unknown training_type {tt!r}")

    @staticmethod
    def _init_simulation_runner(args, device, dataset, model, client_trainer, server_aggregator):
        from .simulation import SimulatorMPI, SimulatorNCCL, SimulatorSingleProcess
        backend = getattr(args, "backend", FEDML_SIMULATION_TYPE_SP)
        if backend in (FEDML_SIMULATION_TYPE_SP, "sp"):
            return SimulatorSingleProcess(args, device, dataset, model, client_trainer, server_aggregator)
        if backend == FEDML_SIMULATION_TYPE_NCCL:
            return SimulatorNCCL(args, device, dataset, model, client_trainer, server_aggregator)
        if backend == FEDML_SIMULATION_TYPE_MPI:
            return SimulatorMPI(args, device, dataset, model, client_trainer, server_aggregator)
        raise ValueError(f"unknown simulation backend {backend!r}")

    @staticmethod
    def _init_cross_silo_runner(args, device, dataset, model, client_trainer, server_aggregator):
        hier = getattr(args, "scenario", "horizontal") == FEDML_CROSS_SILO_SCENARIO_HIERARCHICAL
        role = getattr(args, "role", "server" if int(getattr(args, "rank", 0)) == 0 else "client")
        if hier:
            from .cross_silo.hierarchical import Client, Server
        else:
            from .cross_silo import Client, Server
        if role == "server":
            return Server(args, device, dataset, model, server_aggregator)
        return Client(args, device, dataset, model, client_trainer)

    def run(self):
        return self.runner.run()
