"""Modified Z-score anomaly scorer (C20): ``0.6745 (x - median) / MAD`` over the weighted
degrees of the undirected graph, anomalies where ``|z| > 1``
(All_graphs_IMDB_dataset.ipynb:362-474; recorded output ``['8', '9']`` at :489-491)."""
from __future__ import annotations

from typing import List, Tuple

import numpy as np


def modified_z_scores(values: np.ndarray) -> np.ndarray:
    v = np.asarray(values, dtype=np.float64)
    med = np.median(v)
    mad = np.median(np.abs(v - med))
    if mad == 0:
        return np.zeros_like(v)
    return 0.6745 * (v - med) / mad


def modz_anomalies(values: np.ndarray, threshold: float = 1.0) -> Tuple[List[int], np.ndarray]:
    z = modified_z_scores(values)
    return [int(i) for i in np.flatnonzero(np.abs(z) > threshold)], z
