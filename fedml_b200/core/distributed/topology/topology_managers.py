"""Mixing-matrix generators for gossip / PushSum.

Reference behaviour (symmetric_topology_manager.py:21-56, asymmetric_topology_manager.py:23-81):
ring lattice (Watts-Strogatz k=2, p=0) united with a k-regular ring lattice, unit diagonal,
rows normalised to sum to one; the asymmetric variant additionally flips a fair coin for every
still-missing directed edge (skipping an edge whose reverse was just added).  The reference
builds the lattices through ``networkx.to_numpy_matrix`` which no longer exists in networkx 3
(SURVEY §0.3 item 7); with rewiring probability 0 the lattice is deterministic, so it is built
directly here and no graph library is needed.
"""
import numpy as np

from .base_topology_manager import BaseTopologyManager


def ring_lattice(n: int, k: int) -> np.ndarray:
    """Adjacency of the Watts-Strogatz graph with p=0: node i linked to its k//2 nearest
    neighbours on each side."""
    adj = np.zeros((n, n), dtype=np.float32)
    if n <= 1:
        return adj
    half = max(0, int(k) // 2)
    for i in range(n):
        for d in range(1, half + 1):
            j = (i + d) % n
            if j != i:
                adj[i, j] = adj[j, i] = 1.0
    return adj


def _row_normalise(adj: np.ndarray) -> np.ndarray:
    out = adj.copy()
    deg = (out == 1).sum(axis=1, keepdims=True)
    return out / np.maximum(deg, 1)


class _WeightedTopology(BaseTopologyManager):
    n: int
    topology: np.ndarray

    def get_in_neighbor_weights(self, node_index):
        if node_index >= self.n:
            return []
        return self.topology[:, node_index] if self.directed_columns else self.topology[node_index]

    def get_out_neighbor_weights(self, node_index):
        if node_index >= self.n:
            return []
        return self.topology[node_index]

    def get_in_neighbor_idx_list(self, node_index):
        w = self.get_in_neighbor_weights(node_index)
        return [i for i, v in enumerate(w) if v > 0 and i != node_index]

    def get_out_neighbor_idx_list(self, node_index):
        w = self.get_out_neighbor_weights(node_index)
        return [i for i, v in enumerate(w) if v > 0 and i != node_index]


class SymmetricTopologyManager(_WeightedTopology):
    directed_columns = False

    def __init__(self, n, neighbor_num=2):
        self.n, self.neighbor_num, self.topology = int(n), int(neighbor_num), []

    def generate_topology(self):
        adj = np.maximum(ring_lattice(self.n, 2), ring_lattice(self.n, self.neighbor_num))
        np.fill_diagonal(adj, 1)
        self.topology = _row_normalise(adj)


class AsymmetricTopologyManager(_WeightedTopology):
    # the reference reads in-neighbours from column `node_index` of the row-normalised matrix
    directed_columns = True

    def __init__(self, n, undirected_neighbor_num=3, out_directed_neighbor=3):
        self.n = int(n)
        self.undirected_neighbor_num = int(undirected_neighbor_num)
        self.out_directed_neighbor = int(out_directed_neighbor)
        self.topology = []

    def generate_topology(self):
        adj = np.maximum(ring_lattice(self.n, 2), ring_lattice(self.n, self.undirected_neighbor_num))
        np.fill_diagonal(adj, 1)
        added = set()
        for i in range(self.n):
            zeros = [j for j in range(self.n) if adj[i, j] == 0]
            coins = np.random.randint(2, size=len(zeros))
            for c, j in zip(coins, zeros):
                if c == 1 and (j * self.n + i) not in added:
                    adj[i, j] = 1
                    added.add(i * self.n + j)
        self.topology = _row_normalise(adj)
