"""Client/peer -> GPU placement (1 per GPU at N <= 8; 4 per GPU for the 32-client and 8 per GPU for the 64-peer configs of
BASELINE.json).

``block``: contiguous id blocks, ``ceil(N / world)`` per rank.  ``balanced``: longest-processing-time-first on a per-client
cost (its training-token count): clients are sorted by cost and each goes to the least-loaded rank that still has a free
slot.  With skewed client sizes (Dirichlet split) the block placement leaves the synchronous aggregate waiting for the
slowest rank — 46.7 ms of a 308 ms round in round 1's 32-client run; LPT equalises the per-rank work.  Every rank computes
the same table from the same (deterministic) costs, so no communication is needed.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple


def peers_per_rank(num_peers: int, world: int) -> int:
    return (num_peers + world - 1) // world


def local_peer_ids(num_peers: int, rank: int, world: int) -> List[int]:
    ppr = peers_per_rank(num_peers, world)
    return list(range(rank * ppr, min((rank + 1) * ppr, num_peers)))


def owner_of(peer: int, num_peers: int, world: int) -> Tuple[int, int]:
    """(rank, local slot index) holding ``peer`` under the block placement."""
    ppr = peers_per_rank(num_peers, world)
    return peer // ppr, peer % ppr


class Placement:
    """peer -> (rank, slot) table.  ``Placement.block`` reproduces the functions above."""

    def __init__(self, num_peers: int, world: int, table: Dict[int, Tuple[int, int]], kind: str):
        self.N, self.world, self.P, self.kind = num_peers, world, peers_per_rank(num_peers, world), kind
        self._table = dict(table)
        self._local: List[List[int]] = [[] for _ in range(world)]
        for peer in sorted(self._table, key=lambda p: (self._table[p][0], self._table[p][1])):
            self._local[self._table[peer][0]].append(peer)

    @classmethod
    def block(cls, num_peers: int, world: int) -> "Placement":
        return cls(num_peers, world, {p: owner_of(p, num_peers, world) for p in range(num_peers)}, "block")

    @classmethod
    def balanced(cls, num_peers: int, world: int, cost: Sequence[float]) -> "Placement":
        assert len(cost) == num_peers
        P = peers_per_rank(num_peers, world)
        load, used = [0.0] * world, [0] * world
        table: Dict[int, Tuple[int, int]] = {}
        for peer in sorted(range(num_peers), key=lambda p: (-float(cost[p]), p)):
            r = min((r for r in range(world) if used[r] < P), key=lambda r: (load[r], r))
            table[peer] = (r, used[r])
            used[r] += 1
            load[r] += float(cost[peer])
        return cls(num_peers, world, table, "balanced")

    def owner(self, peer: int) -> Tuple[int, int]:
        return self._table[peer]

    def local_ids(self, rank: int) -> List[int]:
        """Peers of ``rank`` in slot order."""
        return list(self._local[rank])

    def loads(self, cost: Sequence[float]) -> List[float]:
        return [sum(float(cost[p]) for p in self._local[r]) for r in range(self.world)]
