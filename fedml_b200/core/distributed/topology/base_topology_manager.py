"""Topology contract for decentralized FL.

The abstract methods are the reference's interface (topology/base_topology_manager.py:4-23): a topology is generated once and
then queried per node for its in / out neighbours and their mixing weights.  The concrete helpers below are what the
decentralized trainers and the tests use on top of it: the full mixing matrix, a row-stochasticity check (DSGD needs it,
PushSum needs column-stochastic out-weights) and the neighbour pairs as (index, weight) tuples."""
import abc
from typing import List, Tuple

import numpy as np


class BaseTopologyManager(abc.ABC):
    # ---- the reference's contract -------------------------------------------------------------------------------------
    @abc.abstractmethod
    def generate_topology(self):
        """Build the mixing matrix (called once before training)."""

    @abc.abstractmethod
    def get_in_neighbor_idx_list(self, node_index):
        """Indices j with a non-zero weight j -> node_index."""

    @abc.abstractmethod
    def get_out_neighbor_idx_list(self, node_index):
        """Indices j with a non-zero weight node_index -> j."""

    @abc.abstractmethod
    def get_in_neighbor_weights(self, node_index):
        """Row of the mixing matrix that node_index applies to what it receives."""

    @abc.abstractmethod
    def get_out_neighbor_weights(self, node_index):
        """Weights node_index applies to what it sends."""

    # ---- shared helpers -------------------------------------------------------------------------------------------------
    def node_count(self) -> int:
        return int(getattr(self, "n", 0) or len(self.get_in_neighbor_weights(0)))

    def mixing_matrix(self) -> np.ndarray:
        """[n, n] matrix whose row i is ``get_in_neighbor_weights(i)``."""
        n = self.node_count()
        return np.stack([np.asarray(self.get_in_neighbor_weights(i), dtype=np.float64) for i in range(n)])

    def is_row_stochastic(self, atol: float = 1e-8) -> bool:
        w = self.mixing_matrix()
        return bool((w >= -atol).all() and np.allclose(w.sum(axis=1), 1.0, atol=atol))

    def in_neighbors(self, node_index) -> List[Tuple[int, float]]:
        w = self.get_in_neighbor_weights(node_index)
        return [(int(j), float(w[j])) for j in self.get_in_neighbor_idx_list(node_index)]
