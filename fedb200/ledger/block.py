"""Fixed 128-byte block header — identical bytes on host and device so ``block_validate`` (K7)
can re-hash headers on the GPU.

layout (little endian):
  u32 index | u32 round | u32 sender | u32 receiver | u64 num_bytes | u32 chunk_bytes |
  u32 sample_count | u64 timestamp_ns | u8 prev_hash[32] | u8 update_root[32] |
  u32 route_len | u8 route[16] | u32 nonce                                   = 128 bytes
``block_hash = SHA256(header)``; ``receiver = 0xFFFFFFFF`` marks a broadcast / aggregate block.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import List

from .merkle import sha256

HEADER_BYTES = 128
_FMT = "<IIIIQIIQ32s32sI16sI"
assert struct.calcsize(_FMT) == HEADER_BYTES
BROADCAST = 0xFFFFFFFF
GENESIS_PREV = b"\x00" * 32


@dataclass
class BlockHeader:
    index: int
    round: int
    sender: int
    receiver: int
    num_bytes: int
    chunk_bytes: int
    sample_count: int
    timestamp_ns: int
    prev_hash: bytes
    update_root: bytes
    route: List[int] = field(default_factory=list)
    nonce: int = 0

    def pack(self) -> bytes:
        route = bytes(int(r) & 0xFF for r in self.route[:16])
        return struct.pack(_FMT, self.index, self.round, self.sender & 0xFFFFFFFF, self.receiver & 0xFFFFFFFF,
                           self.num_bytes, self.chunk_bytes, self.sample_count, self.timestamp_ns, self.prev_hash,
                           self.update_root, min(len(self.route), 16), route.ljust(16, b"\x00"), self.nonce)

    @classmethod
    def unpack(cls, raw: bytes) -> "BlockHeader":
        (index, rnd, sender, receiver, nbytes, chunk, samples, ts, prev, root, rlen, route, nonce) = struct.unpack(_FMT, raw)
        return cls(index, rnd, sender, receiver, nbytes, chunk, samples, ts, prev, root, list(route[:rlen]), nonce)

    def hash(self) -> bytes:
        return sha256(self.pack())

    def to_json(self) -> dict:
        return dict(index=self.index, round=self.round, sender=self.sender, receiver=self.receiver,
                    num_bytes=self.num_bytes, chunk_bytes=self.chunk_bytes, sample_count=self.sample_count,
                    timestamp_ns=self.timestamp_ns, prev_hash=self.prev_hash.hex(), update_root=self.update_root.hex(),
                    route=self.route, nonce=self.nonce, hash=self.hash().hex())
