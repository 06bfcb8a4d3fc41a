"""``fedml.cross_silo`` — Octopus.  ``Client`` / ``Server`` (args, device, dataset, model, model_trainer=None).run()
(reference: cross_silo/client.py, cross_silo/server.py)."""
from .horizontal import FedML_Horizontal


class Server:
    def __init__(self, args, device, dataset, model, model_trainer=None, server_aggregator=None):
        self.args, self.device, self.dataset, self.model = args, device, dataset, model
        self.model_trainer, self.server_aggregator = model_trainer, server_aggregator
        if hasattr(model_trainer, "aggregate") and server_aggregator is None and model_trainer is not None:
            self.server_aggregator, self.model_trainer = model_trainer, None

    def run(self):
        n = int(getattr(self.args, "client_num_per_round", 1))
        ids = self.args.client_ids() if hasattr(self.args, "client_ids") else []
        return FedML_Horizontal(self.args, 0, len(ids) or n, None, self.device, self.dataset, self.model,
                                self.model_trainer, server_aggregator=self.server_aggregator)


class Client:
    def __init__(self, args, device, dataset, model, model_trainer=None):
        self.args, self.device, self.dataset, self.model, self.model_trainer = args, device, dataset, model, model_trainer

    def run(self):
        n = int(getattr(self.args, "client_num_per_round", 1))
        ids = self.args.client_ids() if hasattr(self.args, "client_ids") else []
        return FedML_Horizontal(self.args, int(self.args.rank), len(ids) or n, None, self.device, self.dataset, self.model,
                                self.model_trainer)


__all__ = ["Client", "Server", "FedML_Horizontal"]


def __getattr__(name):
    """``fedml.cross_silo.hierarchical`` without importing torch.distributed at package import time."""
    if name in ("hierarchical", "horizontal"):
        import importlib
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
