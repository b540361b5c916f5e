"""Mixing-matrix topologies for decentralised training: a ring lattice plus extra regular links, row-normalised,
symmetric or with random directed extra links (reference: ``fedml_core/distributed/topology/symmetric_topology_manager.py:7-78``,
``asymmetric_topology_manager.py:7-103``; those build ``watts_strogatz_graph(n, k, 0)`` with networkx -- rewiring probability 0
is just a ring lattice, generated here directly with numpy)."""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import List

import numpy as np


def ring_lattice(n: int, k: int) -> np.ndarray:
    """Adjacency of the regular ring lattice where each node links to its k//2 nearest neighbours on both sides."""
    a = np.zeros((n, n), dtype=np.float32)
    for i in range(n):
        for d in range(1, k // 2 + 1):
            a[i, (i + d) % n] = 1
            a[i, (i - d) % n] = 1
    return a


class BaseTopologyManager(ABC):
    @abstractmethod
    def generate_topology(self):
        ...

    @abstractmethod
    def get_in_neighbor_idx_list(self, node_index):
        ...

    @abstractmethod
    def get_out_neighbor_idx_list(self, node_index):
        ...

    @abstractmethod
    def get_in_neighbor_weights(self, node_index):
        ...

    @abstractmethod
    def get_out_neighbor_weights(self, node_index):
        ...


class SymmetricTopologyManager(BaseTopologyManager):
    def __init__(self, n: int, neighbor_num: int = 2):
        self.n = n
        self.neighbor_num = neighbor_num
        self.topology = np.zeros((n, n), dtype=np.float32)

    def generate_topology(self):
        t = np.maximum(ring_lattice(self.n, 2), ring_lattice(self.n, int(self.neighbor_num)))
        np.fill_diagonal(t, 1)
        self.topology = t / t.sum(axis=1, keepdims=True)
        return self.topology

    def get_in_neighbor_weights(self, node_index):
        return [] if node_index >= self.n else self.topology[node_index]

    def get_out_neighbor_weights(self, node_index):
        return [] if node_index >= self.n else self.topology[node_index]

    def get_in_neighbor_idx_list(self, node_index) -> List[int]:
        return [i for i, w in enumerate(self.get_in_neighbor_weights(node_index)) if w > 0 and i != node_index]

    def get_out_neighbor_idx_list(self, node_index) -> List[int]:
        return [i for i, w in enumerate(self.get_out_neighbor_weights(node_index)) if w > 0 and i != node_index]


class AsymmetricTopologyManager(BaseTopologyManager):
    def __init__(self, n: int, undirected_neighbor_num: int = 3, out_directed_neighbor: int = 3, seed: int = None):
        self.n = n
        self.undirected_neighbor_num = undirected_neighbor_num
        self.out_directed_neighbor = out_directed_neighbor
        # no seed: numpy's global stream, exactly what the reference draws from; a seed gives a private, reproducible stream
        self.rng = np.random.RandomState(seed) if seed is not None else np.random
        self.topology = np.zeros((n, n), dtype=np.float32)

    def generate_topology(self):
        t = np.maximum(ring_lattice(self.n, 2), ring_lattice(self.n, int(self.undirected_neighbor_num)))
        np.fill_diagonal(t, 1)
        out_links = set()
        for i in range(self.n):
            zeros = [j for j in range(self.n) if t[i, j] == 0]
            pick = self.rng.randint(2, size=len(zeros))
            for sel, j in zip(pick, zeros):
                if sel == 1 and (j * self.n + i) not in out_links:      # never add both directions of a random link
                    t[i, j] = 1
                    out_links.add(i * self.n + j)
        self.topology = t / t.sum(axis=1, keepdims=True)
        return self.topology

    def get_in_neighbor_weights(self, node_index):
        return [] if node_index >= self.n else [self.topology[r][node_index] for r in range(self.n)]

    def get_out_neighbor_weights(self, node_index):
        return [] if node_index >= self.n else self.topology[node_index]

    def get_in_neighbor_idx_list(self, node_index) -> List[int]:
        return [i for i, w in enumerate(self.get_in_neighbor_weights(node_index)) if w > 0 and i != node_index]

    def get_out_neighbor_idx_list(self, node_index) -> List[int]:
        return [i for i, w in enumerate(self.get_out_neighbor_weights(node_index)) if w > 0 and i != node_index]
