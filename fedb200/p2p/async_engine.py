"""Asynchronous gossip engine (C25, K3 host side).

The reference specifies asynchrony only through its timing model — a synchronous exchange
costs the *sum* of the per-destination transfer times, an asynchronous one the *max*
(Medical_Transcriptions_All_graphs.ipynb:979-980, :997-999).  This engine makes it executable:

* every peer owns a monotone epoch counter (never reset), published with its weights;
* a peer mixes whatever neighbour buffers have *landed* by the time it finishes its own local
  epoch — it never waits for a straggler;
* staleness ``tau_j = epoch_i - epoch_j`` (>= 0) discounts a neighbour's mixing weight,
  ``alpha_j = base_j / (1 + tau_j) ** decay``; the self weight absorbs the remainder.  The weights
  are a pure function of the observed epochs, so a CPU replay of the event log reproduces every
  mix exactly (SURVEY §7.4 "async correctness").

Time is virtual (event-driven): compute durations come from measurements or a model, transfer
durations from the path optimiser (``size / bandwidth`` along the chosen route), so the same
engine drives an 8-GPU run and a 64-peer simulation.
"""
from __future__ import annotations

import heapq
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

from .path_opt import optimise_route
from .topology import Topology


def staleness_weight(base: float, tau: int, decay: float = 0.5) -> float:
    return float(base) / float((1 + max(int(tau), 0)) ** decay)


@dataclass(order=True)
class _Event:
    time: float
    seq: int
    kind: str = field(compare=False)          # "trained" | "landed"
    peer: int = field(compare=False)
    src: int = field(compare=False, default=-1)
    epoch: int = field(compare=False, default=0)


@dataclass
class MixRecord:
    time: float
    peer: int
    epoch: int
    sources: List[int]
    source_epochs: List[int]
    weights: List[float]            # aligned with ``sources``; self weight last
    self_weight: float


class AsyncGossipEngine:
    """Event-driven asynchronous gossip over ``n`` peers.

    ``train_fn(peer) -> seconds`` performs one local epoch (and returns its duration, real or
    modelled); ``mix_fn(peer, sources, weights, self_weight)`` applies
    ``w_i <- self_weight * w_i + sum_j weights[j] * inbox[j]``; ``publish_fn(peer, dst)`` copies
    the sender's weights into the destination's inbox slot (the SHA/ledger hook lives there).
    """

    def __init__(self, topo: Topology, nbytes: float, train_fn: Callable[[int], float],
                 mix_fn: Callable[[int, List[int], List[float], float], None],
                 publish_fn: Optional[Callable[[int, int], None]] = None, decay: float = 0.5,
                 fanout: Optional[int] = None, masked: Sequence[int] = (), route_optimizer: bool = True,
                 mix_rate: float = 0.5, min_latency: float = 0.0,
                 land_fn: Optional[Callable[[int, int, int, bool], None]] = None):
        self.topo, self.nbytes, self.decay, self.mix_rate = topo, nbytes, decay, mix_rate
        # visibility lag: a published version is never mixable earlier than this after its publish, whatever the link
        # model says.  On one NVSwitch box the modelled link latencies are ~1 ms against ~100 ms epochs, so without a lag
        # every mix depends on the publish just before it and the real execution serialises across GPUs.
        self.min_latency = float(min_latency)
        self.train_fn, self.mix_fn, self.publish_fn = train_fn, mix_fn, publish_fn
        # land_fn(dst, src, epoch, accepted): the version (src, epoch) has arrived at dst; ``accepted`` is False when a newer
        # one of the same source is already in dst's inbox (or dst is masked).  The data plane keeps one snapshot per
        # published VERSION and binds the inbox entry here, so a source that republishes before its earlier version has
        # landed can never change the bytes a logged mix refers to.
        self.land_fn = land_fn
        self.masked = set(int(m) for m in masked)
        n = topo.n
        self.epoch = np.zeros(n, dtype=np.int64)
        self.clock = np.zeros(n)
        self.inbox: List[Dict[int, int]] = [dict() for _ in range(n)]    # peer -> {src: epoch landed}
        self.log: List[MixRecord] = []
        self.comm_time = np.zeros(n)                                      # async info-passing time per publish
        self.sync_equiv_time = np.zeros(n)
        self.plans = {}
        for i in range(n):
            if i in self.masked:
                continue
            if route_optimizer:
                self.plans[i] = optimise_route(topo, nbytes, i, fanout=fanout, removed=self.masked)
            else:
                hop = topo.hop_seconds(nbytes)[i]
                dst = [j for j in range(n) if j != i and j not in self.masked and np.isfinite(hop[j])]
                dst.sort(key=lambda j: (hop[j], j))
                dst = dst if fanout is None else dst[:fanout]
                from .path_opt import RoutePlan
                lat = {j: float(hop[j]) for j in dst}
                self.plans[i] = RoutePlan(i, {j: [i, j] for j in dst}, lat, max(lat.values(), default=0.0),
                                          sum(lat.values()), max(lat.values(), default=0.0), dst)

    def set_masked(self, masked: Sequence[int]) -> None:
        """PageRank may eliminate peers mid-run; their pending buffers are simply never mixed."""
        self.masked = set(int(m) for m in masked)

    def _weights(self, peer: int) -> Tuple[List[int], List[int], List[float], float]:
        srcs = sorted(s for s in self.inbox[peer] if s not in self.masked)
        if not srcs:
            return [], [], [], 1.0
        base = self.mix_rate / len(srcs)
        eps = [int(self.inbox[peer][s]) for s in srcs]
        ws = [staleness_weight(base, int(self.epoch[peer]) - e, self.decay) for e in eps]
        return srcs, eps, ws, 1.0 - float(sum(ws))

    def run(self, epochs_per_peer: int) -> Dict[str, float]:
        n = self.topo.n
        heap: List[_Event] = []
        seq = 0
        active = [i for i in range(n) if i not in self.masked]
        for i in active:                                   # everyone starts training at t=0
            dur = float(self.train_fn(i))
            seq += 1
            heapq.heappush(heap, _Event(dur, seq, "trained", i, epoch=1))
        done = {i: 0 for i in active}
        while heap:
            ev = heapq.heappop(heap)
            if ev.peer in self.masked:
                if ev.kind == "landed" and self.land_fn is not None:
                    self.land_fn(ev.peer, ev.src, int(ev.epoch), False)
                continue
            if ev.kind == "landed":
                prev = self.inbox[ev.peer].get(ev.src, -1)
                ok = ev.epoch > prev                       # monotone: never replace newer by older
                if ok:
                    self.inbox[ev.peer][ev.src] = ev.epoch
                if self.land_fn is not None:
                    self.land_fn(ev.peer, ev.src, int(ev.epoch), bool(ok))
                continue
            i = ev.peer
            self.epoch[i] = ev.epoch
            self.clock[i] = ev.time
            srcs, eps, ws, sw = self._weights(i)
            self.mix_fn(i, srcs, ws, sw)
            self.log.append(MixRecord(ev.time, i, int(ev.epoch), srcs, eps, ws, sw))
            self.inbox[i].clear()
            plan = self.plans.get(i)
            if plan is not None:
                lat = {j: t for j, t in plan.latency.items() if j not in self.masked}
                for j, t in lat.items():
                    if self.publish_fn is not None:
                        self.publish_fn(i, j)
                    seq += 1
                    heapq.heappush(heap, _Event(ev.time + max(t, self.min_latency), seq, "landed", j, src=i, epoch=int(ev.epoch)))
                self.comm_time[i] += max(lat.values(), default=0.0)
                self.sync_equiv_time[i] += sum(lat.values())
            done[i] += 1
            if done[i] < epochs_per_peer:
                dur = float(self.train_fn(i))
                seq += 1
                heapq.heappush(heap, _Event(ev.time + dur, seq, "trained", i, epoch=int(ev.epoch) + 1))
        return {"makespan": float(self.clock.max()) if len(active) else 0.0,
                "async_comm_time": float(self.comm_time.max()),
                "sync_equiv_comm_time": float(self.sync_equiv_time.max()),
                "mixes": len(self.log)}

    def replay_weights(self) -> List[Tuple[List[float], float]]:
        """Recompute every mix's weights from the logged epochs alone (verification hook)."""
        out = []
        for r in self.log:
            if not r.sources:
                out.append(([], 1.0))
                continue
            base = self.mix_rate / len(r.sources)
            ws = [staleness_weight(base, r.epoch - e, self.decay) for e in r.source_epochs]
            out.append((ws, 1.0 - float(sum(ws))))
        return out
