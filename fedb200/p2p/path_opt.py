"""Information-passing-time model (C22) and weight-transfer path optimiser (C23).

Only formulas exist in the reference:

* per-hop time = model size / channel bandwidth; a multi-hop transfer costs the sum along the
  shortest path (worked example 1->4 via 3, Medical_Transcriptions_All_graphs.ipynb:984-996);
* **synchronous** time = sum over destinations of the path time from the source;
  **asynchronous** = max over destinations (:979-980, :997-999);
* optimiser objective: minimise ``Dg + max_{j in S} latency(i -> j)`` over a selected node
  subset S (All_graphs_IMDB_dataset.ipynb:21, cell 0), ``Dg`` being the local compute time.

The published bar-chart figures cannot be re-derived from the published graph (BASELINE.md
notes), so the formulas are the specification.
"""
from __future__ import annotations

import heapq
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from .topology import Topology


def shortest_paths(cost: np.ndarray, src: int) -> Tuple[np.ndarray, np.ndarray]:
    """Dijkstra on a dense cost matrix (``inf`` = no edge) -> (dist, predecessor)."""
    n = cost.shape[0]
    dist = np.full(n, np.inf)
    pred = np.full(n, -1, dtype=np.int64)
    dist[src] = 0.0
    heap = [(0.0, src)]
    done = np.zeros(n, dtype=bool)
    while heap:
        d, u = heapq.heappop(heap)
        if done[u]:
            continue
        done[u] = True
        for v in range(n):
            c = cost[u, v]
            if v != u and np.isfinite(c) and d + c < dist[v]:
                dist[v] = d + c
                pred[v] = u
                heapq.heappush(heap, (dist[v], v))
    return dist, pred


def path_to(pred: np.ndarray, src: int, dst: int) -> List[int]:
    if dst == src:
        return [src]
    if pred[dst] < 0:
        return []
    out = [dst]
    while out[-1] != src:
        out.append(int(pred[out[-1]]))
    return out[::-1]


def info_passing_time(topo: Topology, nbytes: float, src: int = 0, dests: Optional[Iterable[int]] = None,
                      removed: Iterable[int] = (), routed: bool = True) -> Dict[str, float]:
    """sync = sum, async = max of the per-destination path times from ``src`` (C22)."""
    removed = set(int(r) for r in removed)
    cost = topo.hop_seconds(nbytes)
    for r in removed:
        cost[r, :] = np.inf
        cost[:, r] = np.inf
    if routed:
        dist, _ = shortest_paths(cost, src)
    else:
        dist = cost[src].copy()
        dist[src] = 0.0
    ds = [d for d in (range(topo.n) if dests is None else dests) if d != src and d not in removed]
    times = np.array([dist[d] for d in ds]) if ds else np.zeros(0)
    finite = times[np.isfinite(times)]
    return {"sync": float(finite.sum()), "async": float(finite.max()) if len(finite) else 0.0,
            "unreachable": int(len(times) - len(finite)), "destinations": len(ds)}


@dataclass
class RoutePlan:
    source: int
    routes: Dict[int, List[int]]          # destination -> node path (source ... destination)
    latency: Dict[int, float]             # destination -> seconds
    objective: float                      # Dg + max_j latency
    sync_time: float
    async_time: float
    selected: List[int] = field(default_factory=list)

    def neighbours_in(self) -> List[int]:
        return sorted(self.routes)


def optimise_route(topo: Topology, nbytes: float, src: int, compute_time: float = 0.0,
                   candidates: Optional[Sequence[int]] = None, fanout: Optional[int] = None,
                   removed: Iterable[int] = ()) -> RoutePlan:
    """Choose the subset S (|S| = ``fanout``; all candidates if ``None``) and per-destination
    multi-hop routes that minimise ``Dg + max_{j in S} latency(src -> j)`` (C23).

    With Dijkstra latencies the max-term is minimised by taking the ``fanout`` cheapest
    destinations; multi-hop relays are used whenever they beat the direct channel.
    """
    removed = set(int(r) for r in removed)
    cost = topo.hop_seconds(nbytes)
    for r in removed:
        cost[r, :] = np.inf
        cost[:, r] = np.inf
    dist, pred = shortest_paths(cost, src)
    cand = [c for c in (range(topo.n) if candidates is None else candidates)
            if c != src and c not in removed and np.isfinite(dist[c])]
    cand.sort(key=lambda c: (dist[c], c))
    sel = cand if fanout is None else cand[:fanout]
    routes = {j: path_to(pred, src, j) for j in sel}
    lat = {j: float(dist[j]) for j in sel}
    mx = max(lat.values()) if lat else 0.0
    return RoutePlan(src, routes, lat, compute_time + mx, float(sum(lat.values())), mx, list(sel))
