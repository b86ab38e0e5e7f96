"""Update scorer: pairwise cosine similarity of the flattened client deltas -> transition matrix
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


def similarity_to_transition(cos: np.ndarray) -> np.ndarray:
    """Edge weight u->v = max(cos, 0), no self loops: agreement is endorsement."""
    w = np.clip(np.asarray(cos, dtype=np.float64), 0.0, None)
    np.fill_diagonal(w, 0.0)
    return w


def score_updates(gram: np.ndarray, k_sigma: float = 1.0, low_only: bool = True
                  ) -> Tuple[List[int], np.ndarray, Tuple[float, float]]:
    """PageRank over the similarity graph; peers whose rank falls below ``mean - k·sigma`` (or
    outside the band when ``low_only=False``, the reference's two-sided rule) are anomalous."""
    cos = gram_to_cosine(np.asarray(gram, dtype=np.float64))
    r, _ = pagerank(similarity_to_transition(cos))
    mean = r.sum() / len(r)
    std = float(np.sqrt(((r - mean) ** 2).sum() / len(r)))
    lo, hi = mean - k_sigma * std, mean + k_sigma * std
    bad = (r < lo) if low_only else ((r < lo) | (r > hi))
    if std < 1e-9:
        bad = np.zeros_like(bad)
    return [int(i) for i in np.flatnonzero(bad)], r, (float(lo), float(hi))
