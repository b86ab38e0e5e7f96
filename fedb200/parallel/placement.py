"""Client/peer -> GPU placement: contiguous blocks, ``ceil(N / world)`` per rank (1 per GPU at
N <= 8; 4 per GPU for the 32-client and 8 per GPU for the 64-peer configs of BASELINE.json)."""
from __future__ import annotations

from typing import List, Tuple


def peers_per_rank(num_peers: int, world: int) -> int:
    return (num_peers + world - 1) // world


def local_peer_ids(num_peers: int, rank: int, world: int) -> List[int]:
    ppr = peers_per_rank(num_peers, world)
    return list(range(rank * ppr, min((rank + 1) * ppr, num_peers)))


def owner_of(peer: int, num_peers: int, world: int) -> Tuple[int, int]:
    """(rank, local slot index) holding ``peer``."""
    ppr = peers_per_rank(num_peers, world)
    return peer // ppr, peer % ppr
