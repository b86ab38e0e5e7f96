"""FedAvg strategy with Flower's semantics (C8) and the example-weighted metric mean (C9).

Reference: ``fl.server.strategy.FedAvg(fraction_fit=1.0, fraction_evaluate=1.0,
evaluate_metrics_aggregation_fn=weighted_average)`` (src/Servercase/server_IID_IMDB.py:205-209);
everything else is Flower's default: ``min_fit_clients = min_evaluate_clients =
min_available_clients = 2``, ``accept_failures=True``, no server-side ``evaluate_fn``, no
``initial_parameters`` (initial weights are pulled from one random client).  Aggregation is the
per-tensor ``sum_k n_k w_k / sum_k n_k`` with ``n_k`` whatever ``fit`` returned — the number of
*batches* in the reference (:174).
"""
from __future__ import annotations

import random
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch


def weighted_average(metrics: Sequence[Tuple[int, Dict[str, float]]]) -> Dict[str, float]:
    """Example-weighted mean of ``accuracy`` and ``loss`` (server_IID_IMDB.py:199-203)."""
    examples = [n for n, _ in metrics]
    tot = float(sum(examples)) or 1.0
    keys = set().union(*(m.keys() for _, m in metrics)) if metrics else set()
    return {k: sum(n * m[k] for n, m in metrics if k in m) / tot for k in sorted(keys)}


def aggregate_weighted(results: Sequence[Tuple[Sequence[np.ndarray], int]]) -> List[np.ndarray]:
    """Flower ``aggregate()``: per-tensor weighted mean over ``(ndarrays, num_examples)``."""
    total = float(sum(n for _, n in results))
    out = [np.zeros_like(np.asarray(t), dtype=np.float64) for t in results[0][0]]
    for tensors, n in results:
        for acc, t in zip(out, tensors):
            acc += np.asarray(t, dtype=np.float64) * n
    return [(acc / total).astype(np.float32) for acc in out]


def aggregate_flat(flats: Sequence[torch.Tensor], weights: Sequence[float], out: Optional[torch.Tensor] = None,
                   mask: Optional[Sequence[bool]] = None) -> torch.Tensor:
    """Device-resident version over flat arenas (torch oracle of kernel ``fedavg_reduce``)."""
    w = np.asarray(weights, dtype=np.float64)
    if mask is not None:
        w = w * np.asarray(mask, dtype=np.float64)
    w = w / w.sum()
    acc = torch.zeros_like(flats[0]) if out is None else out.zero_()
    for f, wk in zip(flats, w):
        if wk != 0.0:
            acc.add_(f, alpha=float(wk))
    return acc


class FedAvg:
    def __init__(self, fraction_fit: float = 1.0, fraction_evaluate: float = 1.0, min_fit_clients: int = 2,
                 min_evaluate_clients: int = 2, min_available_clients: int = 2,
                 evaluate_fn: Optional[Callable] = None, initial_parameters=None, accept_failures: bool = True,
                 evaluate_metrics_aggregation_fn: Optional[Callable] = None,
                 fit_metrics_aggregation_fn: Optional[Callable] = None, seed: Optional[int] = None):
        self.fraction_fit, self.fraction_evaluate = fraction_fit, fraction_evaluate
        self.min_fit_clients, self.min_evaluate_clients = min_fit_clients, min_evaluate_clients
        self.min_available_clients = min_available_clients
        self.evaluate_fn, self.initial_parameters = evaluate_fn, initial_parameters
        self.accept_failures = accept_failures
        self.evaluate_metrics_aggregation_fn = evaluate_metrics_aggregation_fn
        self.fit_metrics_aggregation_fn = fit_metrics_aggregation_fn
        self.rng = random.Random(seed)

    # --- sampling ----------------------------------------------------------------------------
    def _sample(self, num_available: int, fraction: float, min_clients: int) -> List[int]:
        if num_available < self.min_available_clients:
            return []
        k = max(int(num_available * fraction), min_clients)
        return sorted(self.rng.sample(range(num_available), min(k, num_available)))

    def configure_fit(self, server_round: int, num_available: int) -> List[int]:
        return self._sample(num_available, self.fraction_fit, self.min_fit_clients)

    def configure_evaluate(self, server_round: int, num_available: int) -> List[int]:
        if self.fraction_evaluate == 0.0:
            return []
        return self._sample(num_available, self.fraction_evaluate, self.min_evaluate_clients)

    # --- aggregation -------------------------------------------------------------------------
    def aggregate_fit(self, server_round: int, results, failures=()):
        """``results``: list of ``(parameters, num_examples, metrics)``; parameters are either a
        list of ndarrays (Flower surface) or one flat tensor (device-resident path)."""
        if not results or (failures and not self.accept_failures):
            return None, {}
        if isinstance(results[0][0], torch.Tensor):
            agg = aggregate_flat([r[0] for r in results], [r[1] for r in results])
        else:
            agg = aggregate_weighted([(r[0], r[1]) for r in results])
        metrics = {}
        if self.fit_metrics_aggregation_fn:
            metrics = self.fit_metrics_aggregation_fn([(r[1], r[2]) for r in results])
        return agg, metrics

    def aggregate_evaluate(self, server_round: int, results, failures=()):
        """``results``: list of ``(loss, num_examples, metrics)`` -> (weighted loss, metrics)."""
        if not results or (failures and not self.accept_failures):
            return None, {}
        tot = float(sum(n for _, n, _ in results)) or 1.0
        loss = sum(l * n for l, n, _ in results) / tot
        metrics = {}
        if self.evaluate_metrics_aggregation_fn:
            metrics = self.evaluate_metrics_aggregation_fn([(n, m) for _, n, m in results])
        return loss, metrics
