from .client_trainer import ClientTrainer
from .server_aggregator import ServerAggregator

__all__ = ["ClientTrainer", "ServerAggregator"]
