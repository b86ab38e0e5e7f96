"""Server simulation driver with Flower's ``start_simulation`` surface (C10).

Reference call: ``fl.simulation.start_simulation(client_fn, num_clients,
config=ServerConfig(num_rounds), strategy, client_resources={"num_cpus":1,"num_gpus":0},
ray_init_args={...,"num_cpus":1})`` (src/Servercase/server_IID_IMDB.py:211-218) — a Ray cluster
pinned to one CPU, so clients run strictly one at a time.  This driver keeps the call shape and
the round structure (init from one random client -> fit round -> aggregate -> evaluate round ->
metric aggregation) without Ray; the multi-GPU runtime (:mod:`fedb200.runtime`) runs the same
strategy with clients spread over ranks and the aggregate done by the fused reduce kernel.
"""
from __future__ import annotations

import random
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

from .fedavg import FedAvg


@dataclass
class ServerConfig:
    num_rounds: int = 1
    round_timeout: Optional[float] = None


@dataclass
class History:
    losses_distributed: List[Tuple[int, float]] = field(default_factory=list)
    losses_centralized: List[Tuple[int, float]] = field(default_factory=list)
    metrics_distributed: Dict[str, List[Tuple[int, float]]] = field(default_factory=dict)
    metrics_distributed_fit: Dict[str, List[Tuple[int, float]]] = field(default_factory=dict)
    metrics_centralized: Dict[str, List[Tuple[int, float]]] = field(default_factory=dict)

    def add_metrics(self, store: Dict[str, List[Tuple[int, float]]], rnd: int, metrics: Dict[str, float]) -> None:
        for k, v in metrics.items():
            store.setdefault(k, []).append((rnd, v))


def start_simulation(client_fn: Callable[[str], object], num_clients: int, config: Optional[ServerConfig] = None,
                     strategy: Optional[FedAvg] = None, client_resources: Optional[dict] = None,
                     ray_init_args: Optional[dict] = None, on_round_end: Optional[Callable] = None,
                     seed: Optional[int] = None) -> History:
    config = config or ServerConfig()
    strategy = strategy or FedAvg()
    hist = History()
    rng = random.Random(seed)
    params = strategy.initial_parameters
    if params is None:                                   # Flower asks one random client
        params = client_fn(str(rng.randrange(num_clients))).get_parameters({})
    if strategy.evaluate_fn is not None:
        res = strategy.evaluate_fn(0, params, {})
        if res is not None:
            hist.losses_centralized.append((0, res[0]))
            hist.add_metrics(hist.metrics_centralized, 0, res[1])
    for rnd in range(1, config.num_rounds + 1):
        results, failures = [], []
        for cid in strategy.configure_fit(rnd, num_clients):
            try:
                results.append(client_fn(str(cid)).fit(params, {}))
            except Exception as exc:                      # accept_failures=True drops the client
                failures.append((cid, exc))
                if not strategy.accept_failures:
                    raise
        agg, fit_metrics = strategy.aggregate_fit(rnd, results, failures)
        if agg is not None:
            params = agg
        hist.add_metrics(hist.metrics_distributed_fit, rnd, fit_metrics)
        if strategy.evaluate_fn is not None:
            res = strategy.evaluate_fn(rnd, params, {})
            if res is not None:
                hist.losses_centralized.append((rnd, res[0]))
                hist.add_metrics(hist.metrics_centralized, rnd, res[1])
        ev, ev_fail = [], []
        for cid in strategy.configure_evaluate(rnd, num_clients):
            try:
                ev.append(client_fn(str(cid)).evaluate(params, {}))
            except Exception as exc:
                ev_fail.append((cid, exc))
                if not strategy.accept_failures:
                    raise
        loss, metrics = strategy.aggregate_evaluate(rnd, ev, ev_fail)
        if loss is not None:
            hist.losses_distributed.append((rnd, loss))
            hist.add_metrics(hist.metrics_distributed, rnd, metrics)
        if on_round_end is not None:
            on_round_end(rnd, params, hist)
    hist.final_parameters = params
    return hist
