"""Community-based anomaly scorer (C21).  The notebook calls it "Louvain" but runs
``nx.community.greedy_modularity_communities`` (Clauset-Newman-Moore) and flags nodes that
belong to no community — which is always empty because CNM partitions every node
(All_graphs_IMDB_dataset.ipynb:548-650; output :665-666).  Dependency-free CNM on a weighted
graph (directed input is symmetrised)."""
from __future__ import annotations

from typing import List, Set

import numpy as np


def greedy_modularity_communities(weights: np.ndarray, weighted: bool = False) -> List[Set[int]]:
    """``weighted=False`` mirrors the notebook's call, which passes no ``weight=`` argument."""
    w = np.asarray(weights, dtype=np.float64)
    if not weighted:
        w = (w > 0).astype(np.float64)
    a = w + w.T
    np.fill_diagonal(a, 0.0)
    n = a.shape[0]
    m2 = a.sum()
    if m2 <= 0:
        return [{i} for i in range(n)]
    comms = {i: {i} for i in range(n)}
    e = {i: {j: a[i, j] / m2 for j in range(n) if a[i, j] > 0} for i in range(n)}   # inter-community edge mass
    deg = {i: a[i].sum() / m2 for i in range(n)}
    while len(comms) > 1:
        best, pair = 0.0, None
        for i in comms:
            for j, eij in e[i].items():
                if j <= i or j not in comms:
                    continue
                dq = 2.0 * (eij - deg[i] * deg[j])
                if dq > best + 1e-15:
                    best, pair = dq, (i, j)
        if pair is None:
            break
        i, j = pair
        comms[i] |= comms.pop(j)
        for k, v in e.pop(j).items():
            if k == i:
                continue
            e[i][k] = e[i].get(k, 0.0) + v
            e[k][i] = e[i][k]
            e[k].pop(j, None)
        e[i].pop(j, None)
        deg[i] += deg.pop(j)
    return sorted(comms.values(), key=lambda s: (-len(s), min(s)))


def community_anomalies(weights: np.ndarray) -> List[int]:
    covered = set().union(*greedy_modularity_communities(weights)) if len(weights) else set()
    return [i for i in range(np.asarray(weights).shape[0]) if i not in covered]
