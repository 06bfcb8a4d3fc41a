"""Server-side operator contract (reference: core/alg_frame/server_aggregator.py:4-35).

The reference's ABC is a copy of ClientTrainer with no aggregation hook.  We keep those
methods and add ``aggregate``/``on_before_aggregation``/``on_after_aggregation`` (the
upstream-FedML extension points) so FedMLRunner can accept a custom server_aggregator.
"""
from abc import ABC, abstractmethod


class ServerAggregator(ABC):
    def __init__(self, model, args=None):
        self.model = model
        self.id = 0
        self.args = args

    def set_id(self, aggregator_id):
        self.id = aggregator_id

    @abstractmethod
    def get_model_params(self):
        ...

    @abstractmethod
    def set_model_params(self, model_parameters):
        ...

    def on_before_aggregation(self, raw_client_model_list):
        return raw_client_model_list

    def aggregate(self, raw_client_model_list):
        """raw_client_model_list: [(num_samples, state_dict)] -> state_dict (sample-weighted mean)."""
        from ..aggregation import fedavg_aggregate
        return fedavg_aggregate(raw_client_model_list)

    def on_after_aggregation(self, aggregated_model):
        return aggregated_model

    def train(self, train_data, device, args):
        return None

    @abstractmethod
    def test(self, test_data, device, args):
        ...

    def test_on_the_server(self, train_data_local_dict, test_data_local_dict, device, args=None) -> bool:
        return False
