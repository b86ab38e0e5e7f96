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


class RelayPlan:
    """Which two-hop routes of one gossip round are executed through a relay GPU, decided identically on every rank.

    ``relays[(i, j)] = r`` says the path optimiser routes source peer ``j`` to destination peer ``i`` through peer ``r``
    (worked example 1 -> 4 via 3, Medical_Transcriptions_All_graphs.ipynb:984-996).  A relay is executed when ``r`` lives on
    a third GPU and that GPU still has a free relay slot: its rank copies ``j``'s slot into ``(rank, slot)`` of the symmetric
    relay buffer (first hop) and ``i``'s row reads that copy (second hop).  One copy of ``j`` on a relay rank serves every
    destination routed through it.  Everything else goes direct."""

    def __init__(self, W: np.ndarray, owner, relays: dict, max_slots: int):
        n = W.shape[0]
        self.copies: dict = {}                 # relay rank -> [(source peer, slot), ...]
        self.via: dict = {}                    # (dst peer, src peer) -> (relay rank, slot)
        self.executed: dict = {}               # (dst peer, src peer) -> executed route [src, (relay,) dst]
        slot_of: dict = {}
        for i in range(n):
            for j in range(n):
                if i == j or W[i, j] == 0.0:
                    continue
                r = relays.get((i, j)) if relays else None
                route = [j, i]
                if r is not None and r not in (i, j):
                    rr = owner(r)[0]
                    if rr != owner(i)[0] and rr != owner(j)[0] and owner(i)[0] != owner(j)[0]:    # a real third GPU
                        key = (j, rr)
                        if key not in slot_of and len(self.copies.get(rr, [])) < max_slots:
                            slot_of[key] = len(self.copies.setdefault(rr, []))
                            self.copies[rr].append((j, slot_of[key]))
                        if key in slot_of:
                            self.via[(i, j)] = (rr, slot_of[key])
                            route = [j, r, i]
                self.executed[(i, j)] = route

    @property
    def any(self) -> bool:
        return bool(self.copies)
