"""Chunked SHA-256 Merkle digest of a byte span (host oracle for the device kernels).

SHA-256 is sequential per message, so one hash over 433 MB cannot use the GPU; the update is cut
into fixed ``chunk_bytes`` leaves hashed independently (``leaf_i = SHA256(chunk_i || 0x00)``, the last
chunk may be short), then reduced pairwise ``SHA256(left || right || 0x01)``.  The one-byte tags separate the two domains
(an interior pair can never be presented as a 64-byte leaf); they are suffixes so the chunk data keeps its 64-byte block
alignment inside the fused copy + hash kernel.  A node without a sibling is promoted unchanged (no duplication).  Same rule
on device, bit-exact against ``hashlib``.
"""
from __future__ import annotations

import hashlib
from typing import List

import numpy as np


LEAF_TAG, NODE_TAG = b"\x00", b"\x01"


def sha256(data: bytes) -> bytes:
    return hashlib.sha256(data).digest()


def leaf_digests(data, chunk_bytes: int) -> List[bytes]:
    mv = memoryview(data).cast("B")
    n = len(mv)
    if n == 0:
        return [sha256(LEAF_TAG)]
    return [sha256(bytes(mv[o:o + chunk_bytes]) + LEAF_TAG) for o in range(0, n, chunk_bytes)]


def reduce_level(nodes: List[bytes]) -> List[bytes]:
    out = [sha256(nodes[i] + nodes[i + 1] + NODE_TAG) for i in range(0, len(nodes) - 1, 2)]
    if len(nodes) % 2:
        out.append(nodes[-1])
    return out


def merkle_root(data, chunk_bytes: int = 4096) -> bytes:
    nodes = leaf_digests(data, chunk_bytes)
    while len(nodes) > 1:
        nodes = reduce_level(nodes)
    return nodes[0]


def root_from_leaves(leaves: np.ndarray) -> bytes:
    """``leaves``: uint8 [n, 32] (e.g. copied back from the device)."""
    nodes = [bytes(r) for r in np.asarray(leaves, dtype=np.uint8).reshape(-1, 32)]
    while len(nodes) > 1:
        nodes = reduce_level(nodes)
    return nodes[0]
