"""Network graph model (C17).

The reference hard-codes a complete directed graph on 10 nodes ``'0'..'9'`` with 90 edges of
``weight = 1/bw`` where ``bw`` is the channel bandwidth in "mbps"
(All_graphs_IMDB_dataset.ipynb:73-167, cell 2; the same 90 literals are pasted into cells 4, 7
and 10).  ``REFERENCE_BANDWIDTH_MBPS[u][v]`` is that table (data, not code), kept as a fixture
so the detectors can be checked against the notebook's recorded outputs.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterable, List, Optional

import numpy as np

REFERENCE_BANDWIDTH_MBPS = np.array([
    [0, 259, 113, 479, 88, 400, 219, 209, 295, 135],
    [252, 0, 145, 343, 247, 421, 303, 383, 387, 272],
    [368, 232, 0, 308, 119, 309, 415, 435, 168, 361],
    [463, 128, 380, 0, 223, 490, 304, 370, 192, 338],
    [401, 479, 402, 465, 0, 285, 291, 370, 447, 205],
    [424, 382, 286, 340, 422, 0, 360, 224, 348, 153],
    [333, 434, 299, 363, 231, 408, 0, 486, 111, 234],
    [243, 426, 188, 180, 489, 192, 415, 0, 378, 148],
    [496, 299, 251, 343, 241, 475, 461, 434, 0, 435],
    [345, 126, 239, 196, 93, 237, 310, 370, 465, 0],
], dtype=np.float64)

NVLINK5_GBPS = 900.0 * 8.0      # 900 GB/s per direction per GPU, in Gbit/s
NVLINK5_MEASURED_GBPS = 770.0 * 8.0


@dataclass
class Topology:
    """Directed bandwidth matrix; ``bw[u, v] == 0`` means no channel u->v."""
    bw: np.ndarray                       # [N, N] bandwidth (any consistent unit)
    unit_bytes_per_s: float = 1e6 / 8.0  # how many bytes/s one unit of ``bw`` is (default: Mbit/s)

    def __post_init__(self):
        self.bw = np.asarray(self.bw, dtype=np.float64).copy()
        np.fill_diagonal(self.bw, 0.0)

    @property
    def n(self) -> int:
        return self.bw.shape[0]

    # ---- constructors -------------------------------------------------------------------
    @classmethod
    def reference10(cls) -> "Topology":
        return cls(REFERENCE_BANDWIDTH_MBPS)

    @classmethod
    def full(cls, n: int, bw: float = NVLINK5_MEASURED_GBPS * 1000.0) -> "Topology":
        """NVSwitch: every peer reaches every other at full link bandwidth."""
        return cls(np.full((n, n), bw))

    @classmethod
    def ring(cls, n: int, bw: float = NVLINK5_MEASURED_GBPS * 1000.0) -> "Topology":
        m = np.zeros((n, n))
        for i in range(n):
            m[i, (i + 1) % n] = bw
            m[i, (i - 1) % n] = bw
        return cls(m)

    @classmethod
    def random(cls, n: int, lo: int = 88, hi: int = 496, seed: int = 0, density: float = 1.0) -> "Topology":
        rng = np.random.default_rng(seed)
        m = rng.integers(lo, hi + 1, size=(n, n)).astype(np.float64)
        if density < 1.0:
            keep = rng.random((n, n)) < density
            for i in range(n):                      # keep a ring so the graph stays connected
                keep[i, (i + 1) % n] = True
            m *= keep
        return cls(m)

    @classmethod
    def tiled(cls, base: "Topology", n: int) -> "Topology":
        """Extend a small measured graph to ``n`` peers by tiling it (64-peer scale-up config)."""
        reps = (n + base.n - 1) // base.n
        return cls(np.tile(base.bw, (reps, reps))[:n, :n], base.unit_bytes_per_s)

    @classmethod
    def from_name(cls, name: str, n: int, seed: int = 0) -> "Topology":
        if name == "full":
            return cls.full(n)
        if name == "ring":
            return cls.ring(n)
        if name == "reference10":
            base = cls.reference10()
            return base if n == 10 else (base.subgraph(range(n)) if n < 10 else cls.tiled(base, n))
        if name == "random":
            return cls.random(n, seed=seed)
        raise ValueError(f"unknown topology {name!r}")

    # ---- views ----------------------------------------------------------------------------
    def edge_weights(self) -> np.ndarray:
        """The notebook's edge attribute ``weight = 1/bw`` (0 where no edge)."""
        with np.errstate(divide="ignore"):
            w = np.where(self.bw > 0, 1.0 / self.bw, 0.0)
        return w

    def adjacency(self) -> np.ndarray:
        return (self.bw > 0).astype(np.float64)

    def hop_seconds(self, nbytes: float) -> np.ndarray:
        """Per-hop time ``size / bandwidth`` (Medical_Transcriptions_All_graphs.ipynb:984-986)."""
        with np.errstate(divide="ignore"):
            return np.where(self.bw > 0, nbytes / (self.bw * self.unit_bytes_per_s), np.inf)

    def weighted_degree_undirected(self) -> np.ndarray:
        """Weighted degree on ``nx.Graph`` built by inserting the directed edges in row-major
        order, so the later duplicate ``(v,u)`` overwrites ``(u,v)`` — the behaviour of cells
        4/7 (All_graphs_IMDB_dataset.ipynb:206-316)."""
        w = self.edge_weights()
        n = self.n
        sym = np.zeros((n, n))
        for u in range(n):
            for v in range(n):
                if u != v and w[u, v] > 0:
                    sym[u, v] = sym[v, u] = w[u, v]
        return sym.sum(axis=1)

    def subgraph(self, keep: Iterable[int]) -> "Topology":
        idx = np.asarray(list(keep), dtype=np.int64)
        return Topology(self.bw[np.ix_(idx, idx)], self.unit_bytes_per_s)

    def without(self, removed: Iterable[int]) -> "Topology":
        r = set(int(x) for x in removed)
        return self.subgraph([i for i in range(self.n) if i not in r])

    def slow_links(self, nodes: Iterable[int], factor: float = 0.05) -> "Topology":
        """Fault injection: scale every channel touching ``nodes`` by ``factor``."""
        bw = self.bw.copy()
        for i in nodes:
            bw[i, :] *= factor
            bw[:, i] *= factor
        return Topology(bw, self.unit_bytes_per_s)
