"""Deterministic asynchronous-gossip schedule (C25 / K3 host side) for multi-GPU execution.

The event engine (:mod:`fedb200.p2p.async_engine`) decides, in virtual time, WHICH published version of each
neighbour a peer mixes and with which staleness weights.  When peers live on different GPUs every rank must
agree on that outcome without talking, so the virtual clock is driven by a *model* (per-peer epoch duration,
``size / bandwidth`` link latencies along the optimised route) instead of measurements: every rank computes the
identical schedule up front and then executes only its own peers' events.  Cross-GPU ordering is enforced on the
device by monotone flag words (``pub[p]`` = last epoch peer p published, ``done[p]`` = last epoch it finished
mixing), never by host barriers — a peer that is ahead keeps going.

Published versions live in a small ring of outbox slots per peer.  Because the whole schedule is known, slot reuse
is computed exactly: version (p, e) may overwrite an older version only after that version's last consumer event,
and the executor waits on precisely those consumers' ``done`` flags.  All waits point to earlier events of one
global order, so the protocol cannot deadlock.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from .async_engine import AsyncGossipEngine
from .topology import Topology


@dataclass
class ScheduledEvent:
    index: int
    time: float
    peer: int
    epoch: int                               # 1-based epoch this peer has just trained
    sources: List[int]
    source_epochs: List[int]
    weights: List[float]
    self_weight: float
    slot: int = 0                            # outbox slot the post-mix state is published to
    overwrite_waits: List[Tuple[int, int]] = field(default_factory=list)   # (consumer peer, consumer epoch) to wait for
    dests: List[int] = field(default_factory=list)
    route_latency: float = 0.0               # async information-passing time of this publish (max over dests)
    route_sum: float = 0.0                   # synchronous equivalent (sum over dests)


@dataclass
class AsyncSchedule:
    events: List[ScheduledEvent]
    slots_per_peer: int
    makespan: float
    async_comm_time: float
    sync_equiv_comm_time: float
    masked: List[int]

    def slot_of(self, peer: int, epoch: int) -> int:
        return self._slot[(peer, epoch)]


def model_durations(n: int, base_s: float = 0.05, spread: float = 0.5, seed: int = 0) -> np.ndarray:
    """Per-peer local-epoch duration model: ``base * U(1 - spread/2, 1 + spread)`` (heterogeneous peers)."""
    rng = np.random.default_rng(seed + 4242)
    return base_s * rng.uniform(1.0 - spread / 2.0, 1.0 + spread, size=n)


def build_async_schedule(topo: Topology, nbytes: float, epochs_per_peer: int, durations: Sequence[float],
                         decay: float = 0.5, mix_rate: float = 0.5, fanout: Optional[int] = None,
                         masked: Sequence[int] = (), route_optimizer: bool = True, max_slots: int = 8,
                         min_latency: float = 0.0) -> AsyncSchedule:
    dur = np.asarray(durations, dtype=np.float64)
    eng = AsyncGossipEngine(topo, nbytes, lambda k: float(dur[k]), lambda *a: None, None, decay, fanout, masked,
                            route_optimizer, mix_rate, min_latency)
    res = eng.run(epochs_per_peer)
    events = []
    for i, r in enumerate(eng.log):
        plan = eng.plans.get(r.peer)
        lat = {j: t for j, t in (plan.latency.items() if plan else []) if j not in eng.masked}
        events.append(ScheduledEvent(i, r.time, r.peer, r.epoch, list(r.sources), list(r.source_epochs), list(r.weights),
                                     r.self_weight, dests=sorted(lat), route_latency=max(lat.values(), default=0.0),
                                     route_sum=sum(lat.values())))
    # ---- exact slot allocation: last consumer of every published version -> greedy ring reuse ------------------
    last_use: Dict[Tuple[int, int], int] = {}
    consumers: Dict[Tuple[int, int], List[Tuple[int, int, int]]] = {}
    for ev in events:
        for s, e in zip(ev.sources, ev.source_epochs):
            last_use[(s, e)] = ev.index
            consumers.setdefault((s, e), []).append((ev.peer, ev.epoch, ev.index))
    slot_of: Dict[Tuple[int, int], int] = {}
    free_at: Dict[int, List[Tuple[int, Tuple[int, int]]]] = {}          # peer -> per-slot (last consumer index, version)
    n_slots = 1
    for ev in events:
        ring = free_at.setdefault(ev.peer, [])
        chosen = None
        for si, (busy_until, _) in enumerate(ring):
            if busy_until < ev.index:                       # every consumer of the old version precedes this publish
                chosen = si
                break
        if chosen is None:
            ring.append((-1, (ev.peer, 0)))
            chosen = len(ring) - 1
            if len(ring) > max_slots:
                raise RuntimeError(f"peer {ev.peer} needs more than {max_slots} outbox versions in flight: link latency is "
                                   "too large relative to the epoch time for on-device buffering")
        old_version = ring[chosen][1]
        ev.slot = chosen
        ev.overwrite_waits = [(c, ce) for c, ce, _ in consumers.get(old_version, [])]
        ring[chosen] = (last_use.get((ev.peer, ev.epoch), ev.index), (ev.peer, ev.epoch))
        slot_of[(ev.peer, ev.epoch)] = chosen
        n_slots = max(n_slots, len(ring))
    sched = AsyncSchedule(events, n_slots, res["makespan"], res["async_comm_time"], res["sync_equiv_comm_time"],
                          sorted(eng.masked))
    sched._slot = slot_of
    return sched


def verify_schedule(sched: AsyncSchedule, num_peers: int, world: int, trials: int = 16, seed: int = 0) -> Dict[str, int]:
    """Protocol-level check of the flag handshake the executor (``FedRuntime.run_async_scheduled``) performs.

    (A) *Ordering*: every wait of an event — ``pub[s] >= e`` for each mixed source version, ``done[c] >= ce`` for each
    consumer of the version its outbox slot still holds — targets an event with a smaller global index, and each rank
    runs its own events in index order; by induction no rank can wait forever.
    (B) *Safety under any interleaving*: ranks are simulated as independent state machines that advance in a random
    order (``trials`` different interleavings), blocking exactly where the executor blocks.  Whenever a mix reads version
    ``(s, e)`` the simulated outbox slot must hold exactly that version (published, and not yet overwritten); a violation
    means a missing or insufficient wait.  Returns counters; raises ``AssertionError`` with the offending event otherwise.
    """
    from ..parallel.placement import owner_of
    by_version = {(ev.peer, ev.epoch): ev for ev in sched.events}
    # ---- (A) every wait points backwards in the one global order ----
    n_waits = 0
    for ev in sched.events:
        for s, e in zip(ev.sources, ev.source_epochs):
            src = by_version.get((s, e))
            assert src is not None, f"event {ev.index} mixes version {(s, e)} that nobody publishes"
            assert src.index < ev.index, f"event {ev.index} waits on pub{(s, e)} published later (event {src.index})"
            n_waits += 1
        for c, ce in ev.overwrite_waits:
            con = by_version.get((c, ce))
            assert con is not None and con.index < ev.index, f"event {ev.index} waits on done{(c, ce)} that is not earlier"
            n_waits += 1
    per_rank: List[List[ScheduledEvent]] = [[] for _ in range(world)]
    for ev in sched.events:
        per_rank[owner_of(ev.peer, num_peers, world)[0]].append(ev)
    for evs in per_rank:
        assert all(a.index < b.index for a, b in zip(evs, evs[1:])), "a rank's events are not in global order"
    # ---- (B) randomised interleavings of the per-rank state machines ----
    rng = np.random.default_rng(seed)
    reads = 0
    for _ in range(trials):
        pub = [0] * num_peers
        done = [0] * num_peers
        slots: Dict[Tuple[int, int], Tuple[int, int]] = {}        # (peer, slot) -> version held
        pc = [[0, 0] for _ in range(world)]                       # per rank: [event cursor, phase 0 = mix, 1 = publish]
        remaining = sum(len(e) for e in per_rank)
        while remaining:
            runnable = []
            for r in range(world):
                i, ph = pc[r]
                if i >= len(per_rank[r]):
                    continue
                ev = per_rank[r][i]
                if ph == 0:
                    ok = all(pub[s] >= e for s, e in zip(ev.sources, ev.source_epochs))
                else:
                    ok = all(done[c] >= ce for c, ce in ev.overwrite_waits)
                if ok:
                    runnable.append(r)
            assert runnable, f"deadlock: cursors {pc}"
            r = int(runnable[rng.integers(len(runnable))])
            i, ph = pc[r]
            ev = per_rank[r][i]
            if ph == 0:
                for s, e in zip(ev.sources, ev.source_epochs):
                    held = slots.get((s, sched.slot_of(s, e)))
                    assert held == (s, e), (f"event {ev.index} (peer {ev.peer} epoch {ev.epoch}) reads version {(s, e)} "
                                            f"but slot {sched.slot_of(s, e)} of peer {s} holds {held}")
                    reads += 1
                done[ev.peer] = ev.epoch
                pc[r][1] = 1
            else:
                slots[(ev.peer, ev.slot)] = (ev.peer, ev.epoch)
                pub[ev.peer] = ev.epoch
                pc[r] = [i + 1, 0]
                remaining -= 1
    return {"events": len(sched.events), "waits": n_waits, "reads_checked": reads, "interleavings": trials}
