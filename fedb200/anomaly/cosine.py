"""Update scorer: pairwise (norm-aware) cosine similarity of the flattened client deltas -> transition matrix
-> PageRank -> ``mean ± k·sigma`` mask.  North-star extension (BASELINE.json): the reference only
scores a latency graph offline (C18); here the same PageRank rule is applied on-line to a
similarity graph of the weight updates so poisoned peers drop out of the aggregate.  This file is
the host oracle for kernels K4 (``cosine_gram``) and K5 (``pagerank_power``)."""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np
import torch

from .pagerank import pagerank


def cosine_gram(deltas: Sequence[torch.Tensor]) -> Tuple[np.ndarray, np.ndarray]:
    """(Gram matrix ``G[i,j] = <d_i, d_j>``, cosine matrix) in fp64 on host."""
    d = torch.stack([x.reshape(-1).to(torch.float64).cpu() for x in deltas])
    g = (d @ d.T).numpy()
    return g, gram_to_cosine(g)


def gram_to_cosine(g: np.ndarray) -> np.ndarray:
    nrm = np.sqrt(np.clip(np.diag(g), 1e-30, None))
    return g / nrm[:, None] / nrm[None, :]


def gram_to_edge_weights(g: np.ndarray) -> np.ndarray:
    """Endorsement weight between two updates from their Gram entries:

        w[u, v] = (1 + cos(u, v)) / 2  *  min(|d_u|, |d_v|) / max(|d_u|, |d_v|),   w[u, u] = 0.

    The direction term keeps the graph connected (a hard clip at 0 would split colluding attackers into
    their own component, where uniform teleport hands them an average rank) and the norm-ratio term sees what
    plain cosine is blind to — a scaled update (same direction, 10x the step).  A random walk on this graph
    concentrates on the majority cluster of mutually consistent, similarly sized updates."""
    g = np.asarray(g, dtype=np.float64)
    d = np.clip(np.diag(g), 1e-30, None)
    cos = np.clip(g / np.sqrt(d[:, None] * d[None, :]), -1.0, 1.0)
    ratio = np.sqrt(np.minimum(d[:, None], d[None, :]) / np.maximum(d[:, None], d[None, :]))
    w = 0.5 * (1.0 + cos) * ratio
    np.fill_diagonal(w, 0.0)
    return w


def similarity_to_transition(gram: np.ndarray) -> np.ndarray:      # kept as the public name of the mapping
    return gram_to_edge_weights(gram)


def score_updates(gram: np.ndarray, k_sigma: float = 1.0, low_only: bool = True, min_rel_gap: float = 0.1
                  ) -> Tuple[List[int], np.ndarray, Tuple[float, float]]:
    """PageRank over the similarity graph; peers whose rank falls below ``mean - k·sigma`` (or
    outside the band when ``low_only=False``, the reference's two-sided rule) are anomalous."""
    r, _ = pagerank(gram_to_edge_weights(gram))
    mean = r.sum() / len(r)
    std = float(np.sqrt(((r - mean) ** 2).sum() / len(r)))
    lo, hi = mean - k_sigma * std, mean + k_sigma * std
    bad = (r < lo) if low_only else ((r < lo) | (r > hi))
    # the reference's mean ± 1 sigma rule always flags ~1/6 of any population; on an all-honest round the ranks are
    # nearly uniform, so additionally require a rank at least `min_rel_gap` below the mean before eliminating a peer
    bad &= np.abs(r - mean) > min_rel_gap * mean
    if std < 1e-9:
        bad = np.zeros_like(bad)
    return [int(i) for i in np.flatnonzero(bad)], r, (float(lo), float(hi))
