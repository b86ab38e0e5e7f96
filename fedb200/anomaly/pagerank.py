"""PageRank anomaly scorer (C18) — CPU oracle for the on-device kernel ``pagerank_power``.

Reference: ``nx.pagerank(G, weight='weight')`` then ``mean ± 1·std`` (population std) two-sided
threshold (All_graphs_IMDB_dataset.ipynb:168-176).  networkx defaults: alpha 0.85, L1 tolerance
``N * 1e-6``, <= 100 iterations, out-weight row normalisation, uniform teleport and uniform
redistribution of dangling mass.  The recorded output (:52-54) is the test oracle:
thresholds (0.083493, 0.116507), anomalies {0, 4, 7, 9}.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np


def pagerank(weights: np.ndarray, alpha: float = 0.85, tol: float = 1e-6, max_iter: int = 100
             ) -> Tuple[np.ndarray, int]:
    """Power iteration on a dense weighted digraph ``weights[u, v]`` (u -> v). Returns (ranks, iters)."""
    w = np.asarray(weights, dtype=np.float64)
    n = w.shape[0]
    out = w.sum(axis=1)
    dangling = out == 0
    p = np.divide(w, out[:, None], out=np.zeros_like(w), where=out[:, None] > 0)
    x = np.full(n, 1.0 / n)
    tele = np.full(n, 1.0 / n)
    for it in range(1, max_iter + 1):
        xl = x
        x = alpha * (xl @ p + xl[dangling].sum() * tele) + (1.0 - alpha) * tele
        if np.abs(x - xl).sum() < n * tol:
            return x, it
    return x, max_iter


def pagerank_anomalies(weights: np.ndarray, k_sigma: float = 1.0, **kw) -> Tuple[List[int], np.ndarray, Tuple[float, float]]:
    r, _ = pagerank(weights, **kw)
    mean = r.sum() / len(r)
    std = float(np.sqrt(((r - mean) ** 2).sum() / len(r)))
    lo, hi = mean - k_sigma * std, mean + k_sigma * std
    return [int(i) for i in np.flatnonzero((r < lo) | (r > hi))], r, (float(lo), float(hi))
