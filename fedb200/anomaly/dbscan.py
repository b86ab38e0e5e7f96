"""DBSCAN anomaly scorer (C19): ``DBSCAN(eps=300, min_samples=2)`` on the 1-D weighted-degree
feature; anomalies are the noise points (label -1).  With the reference's graph the feature scale
is ~1e-2, so eps=300 puts everything in one cluster and the result is empty
(All_graphs_IMDB_dataset.ipynb:206-316; output :331-333).  Dependency-free 1-D implementation
with sklearn's semantics (a point counts itself towards ``min_samples``)."""
from __future__ import annotations

from typing import List, Tuple

import numpy as np


def dbscan_1d(x: np.ndarray, eps: float, min_samples: int) -> np.ndarray:
    x = np.asarray(x, dtype=np.float64).reshape(-1)
    n = len(x)
    nbrs = [np.flatnonzero(np.abs(x - x[i]) <= eps) for i in range(n)]
    core = np.array([len(nb) >= min_samples for nb in nbrs])
    labels = np.full(n, -1, dtype=np.int64)
    cid = 0
    for i in range(n):
        if labels[i] != -1 or not core[i]:
            continue
        stack = [i]
        labels[i] = cid
        while stack:
            p = stack.pop()
            if not core[p]:
                continue
            for q in nbrs[p]:
                if labels[q] == -1:
                    labels[q] = cid
                    stack.append(int(q))
        cid += 1
    return labels


def dbscan_anomalies(values: np.ndarray, eps: float = 300.0, min_samples: int = 2) -> Tuple[List[int], np.ndarray]:
    labels = dbscan_1d(values, eps, min_samples)
    return [int(i) for i in np.flatnonzero(labels == -1)], labels
