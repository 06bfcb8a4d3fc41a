"""Client-side operator contract (reference: core/alg_frame/client_trainer.py:4-40).

A ``ClientTrainer`` is a stateless wrapper around a model: the runner hands it global
parameters, a data shard and ``args``; it returns updated parameters.  The same five
abstract methods as the reference are kept so user subclasses drop in unchanged; the
optional ``get_flat_params``/``train_batched`` hooks are what the B200 runtime uses to keep
parameters device-resident (no per-client ``.cpu()`` round-trip).
"""
from abc import ABC, abstractmethod


class ClientTrainer(ABC):
    def __init__(self, model, args=None):
        self.model = model
        self.id = 0
        self.args = args

    def set_id(self, trainer_id):
        self.id = trainer_id

    @abstractmethod
    def get_model_params(self):
        ...

    @abstractmethod
    def set_model_params(self, model_parameters):
        ...

    @abstractmethod
    def train(self, train_data, device, args):
        ...

    @abstractmethod
    def test(self, test_data, device, args):
        ...

    def test_on_the_server(self, train_data_local_dict, test_data_local_dict, device, args=None) -> bool:
        """Return True when the trainer evaluated by itself and the default eval must be skipped."""
        return False
