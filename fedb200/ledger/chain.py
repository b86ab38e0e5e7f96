"""Append-only hash chain + validation with tamper detection (C24).

Consensus is deliberately minimal (the reference specifies none): every participant appends the
same deterministic blocks and any participant can validate the whole chain — on the host here, or
on the device through ``block_validate`` — and check a received update against its recorded
Merkle root before mixing it.  An optional proof-of-work difficulty (leading zero bits) exists for
experiments on ledger cost; it defaults to 0.
"""
from __future__ import annotations

import json
import time
from typing import Iterable, List, Optional, Sequence

import numpy as np

from .block import BROADCAST, GENESIS_PREV, HEADER_BYTES, BlockHeader
from .merkle import merkle_root


class LedgerError(RuntimeError):
    pass


def _leading_zero_bits(h: bytes) -> int:
    n = 0
    for b in h:
        if b == 0:
            n += 8
            continue
        return n + (8 - b.bit_length())
    return n


class Ledger:
    def __init__(self, chunk_bytes: int = 4096, difficulty_bits: int = 0, clock=time.time_ns):
        self.chunk_bytes, self.difficulty_bits, self.clock = chunk_bytes, difficulty_bits, clock
        self.blocks: List[BlockHeader] = []
        self.hashes: List[bytes] = []

    @property
    def height(self) -> int:
        return len(self.blocks)

    @property
    def head(self) -> bytes:
        return self.hashes[-1] if self.hashes else GENESIS_PREV

    def append(self, round_idx: int, sender: int, receiver: int, update_root: bytes, num_bytes: int,
               sample_count: int = 0, route: Sequence[int] = (), timestamp_ns: Optional[int] = None) -> BlockHeader:
        blk = BlockHeader(self.height, round_idx, sender, receiver, num_bytes, self.chunk_bytes, sample_count,
                          self.clock() if timestamp_ns is None else timestamp_ns, self.head, bytes(update_root),
                          list(route))
        while _leading_zero_bits(blk.hash()) < self.difficulty_bits:
            blk.nonce += 1
        self.blocks.append(blk)
        self.hashes.append(blk.hash())
        return blk

    def record_update(self, round_idx: int, sender: int, receiver: int, payload, sample_count: int = 0,
                      route: Sequence[int] = (), timestamp_ns: Optional[int] = None) -> BlockHeader:
        """Hash ``payload`` (bytes-like) on the host and append its block."""
        mv = memoryview(payload).cast("B")
        return self.append(round_idx, sender, receiver, merkle_root(mv, self.chunk_bytes), len(mv), sample_count,
                           route, timestamp_ns)

    def verify_update(self, index: int, payload) -> bool:
        mv = memoryview(payload).cast("B")
        b = self.blocks[index]
        return len(mv) == b.num_bytes and merkle_root(mv, b.chunk_bytes) == b.update_root

    def validate(self) -> int:
        """Returns -1 if the chain is intact, else the index of the first bad block."""
        prev = GENESIS_PREV
        for i, b in enumerate(self.blocks):
            h = b.hash()
            if b.index != i or b.prev_hash != prev or h != self.hashes[i] \
                    or _leading_zero_bits(h) < self.difficulty_bits:
                return i
            prev = h
        return -1

    def assert_valid(self) -> None:
        bad = self.validate()
        if bad >= 0:
            raise LedgerError(f"ledger corrupt at block {bad}")

    # ---- serialisation (binary headers are what the device validator consumes) -----------
    def headers_array(self) -> np.ndarray:
        raw = b"".join(b.pack() for b in self.blocks)
        return np.frombuffer(raw, dtype=np.uint8).reshape(-1, HEADER_BYTES).copy()

    def to_json(self) -> str:
        return json.dumps({"chunk_bytes": self.chunk_bytes, "difficulty_bits": self.difficulty_bits,
                           "blocks": [b.to_json() for b in self.blocks]})

    def save(self, path: str) -> None:
        with open(path, "w") as fh:
            fh.write(self.to_json())

    @classmethod
    def load(cls, path: str) -> "Ledger":
        with open(path) as fh:
            d = json.load(fh)
        led = cls(d["chunk_bytes"], d.get("difficulty_bits", 0))
        for j in d["blocks"]:
            b = BlockHeader(j["index"], j["round"], j["sender"], j["receiver"], j["num_bytes"], j["chunk_bytes"],
                            j["sample_count"], j["timestamp_ns"], bytes.fromhex(j["prev_hash"]),
                            bytes.fromhex(j["update_root"]), j["route"], j["nonce"])
            led.blocks.append(b)
            led.hashes.append(bytes.fromhex(j["hash"]))
        return led
