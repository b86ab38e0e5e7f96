"""Gossip mixing (C11/C12 and kernel K2's host-side contract).

The reference's "fully decentralised" aggregator is the unweighted mean
``[sum(p)/len(p) for p in zip(*aggregated_params)]`` followed by ``load_state_dict``
(src/Serverlesscase/serverless_IID_IMDB.py:269-270) — i.e. a complete-graph mixing matrix with
uniform weights.  General form: ``w_i <- sum_j W[i,j] w_j`` with a row-stochastic ``W`` built
from the topology (uniform over in-neighbours + self, or Metropolis-Hastings); anomalous peers
are masked out column-wise and rows are renormalised.
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Sequence

import numpy as np
import torch


def mixing_matrix(adj: np.ndarray, kind: str = "uniform") -> np.ndarray:
    """Row-stochastic mixing matrix from an adjacency (``adj[i, j] > 0``: i hears from j)."""
    a = (np.asarray(adj) > 0).astype(np.float64)
    n = a.shape[0]
    np.fill_diagonal(a, 1.0)
    if kind == "uniform":
        return a / a.sum(axis=1, keepdims=True)
    if kind == "metropolis":
        deg = a.sum(axis=1) - 1.0
        w = np.zeros((n, n))
        for i in range(n):
            for j in range(n):
                if i != j and a[i, j] > 0:
                    w[i, j] = 1.0 / (1.0 + max(deg[i], deg[j]))
            w[i, i] = 1.0 - w[i].sum()
        return w
    raise ValueError(kind)


def apply_mask(w: np.ndarray, masked: Iterable[int]) -> np.ndarray:
    """Drop masked peers' contributions and renormalise each row (masked rows keep themselves)."""
    w = np.asarray(w, dtype=np.float64).copy()
    m = sorted(set(int(x) for x in masked))
    if not m:
        return w
    w[:, m] = 0.0
    for i in range(w.shape[0]):
        s = w[i].sum()
        if s <= 0.0:
            w[i, i] = 1.0
        else:
            w[i] /= s
    return w


def mix_flat(flats: Sequence[torch.Tensor], weights: Sequence[float], out: Optional[torch.Tensor] = None
             ) -> torch.Tensor:
    """``out = sum_j weights[j] * flats[j]`` in fp32 — host/torch oracle for ``gossip_mix``."""
    acc = torch.zeros_like(flats[0], dtype=torch.float32) if out is None else out.zero_()
    for f, w in zip(flats, weights):
        if w != 0.0:
            acc.add_(f.to(torch.float32), alpha=float(w))
    return acc
