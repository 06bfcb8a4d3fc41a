"""Topology contract for decentralized FL (reference: topology/base_topology_manager.py:4-23)."""
import abc


class BaseTopologyManager(abc.ABC):
    @abc.abstractmethod
    def generate_topology(self):
        ...

    @abc.abstractmethod
    def get_in_neighbor_idx_list(self, node_index):
        ...

    @abc.abstractmethod
    def get_out_neighbor_idx_list(self, node_index):
        ...

    @abc.abstractmethod
    def get_in_neighbor_weights(self, node_index):
        ...

    @abc.abstractmethod
    def get_out_neighbor_weights(self, node_index):
        ...
