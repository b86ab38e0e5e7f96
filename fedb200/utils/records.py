"""Per-round peer records exchanged on the device.

Every round each rank has a few numbers per local peer that every other rank needs: the Merkle root of the published
update (ledger block), ``len(trainloader)`` (the FedAvg weight, src/Servercase/server_IID_IMDB.py:174), the last
training loss and the evaluation sums (``weighted_average``, :199-203; ``local_accuracy`` print,
src/Serverlesscase/serverless_IID_IMDB.py:265).  Round 1 moved them with a pickled ``dist.all_gather_object`` — a host
collective in every round.  Here they live in one symmetric ``[world, P * REC_WORDS]`` int32 table: a rank fills the rows
of its own peers with device-side copies (results never visit the host), ``Comm.exchange_slots`` pushes the row block
to every peer inside one small kernel, and the whole table comes back with the round's single ``HostReadback.sync``.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np
import torch

REC_WORDS = 24
# word offsets inside one peer record
R_ROOT, R_PEER, R_BATCHES, R_SAMPLES, R_LOSS, R_LCORRECT, R_LTOTAL = 0, 8, 9, 10, 11, 12, 13
R_ELOSS, R_ECORRECT, R_ENDS, R_ETOTAL, R_EBATCHES, R_VALID = 14, 15, 16, 17, 18, 19


class RoundRecords:
    def __init__(self, comm, world: int, rank: int, peers_per_rank: int, device):
        self.comm, self.world, self.rank, self.P = comm, world, rank, peers_per_rank
        self.device = torch.device(device)
        self.table = comm.alloc("round_records", world * peers_per_rank * REC_WORDS, torch.int32).view(
            world, peers_per_rank * REC_WORDS)
        self._host = torch.zeros(peers_per_rank, REC_WORDS, dtype=torch.int32,
                                 pin_memory=self.device.type == "cuda")

    def _rec(self, li: int) -> torch.Tensor:
        return self.table[self.rank, li * REC_WORDS:(li + 1) * REC_WORDS]

    def begin(self, infos: List[Dict]) -> None:
        """Host-known fields of this rank's peers (one pinned H2D copy) — call before the device-side setters."""
        h = self._host
        h.zero_()
        for li, i in enumerate(infos):
            h[li, R_PEER], h[li, R_BATCHES], h[li, R_SAMPLES], h[li, R_VALID] = i["peer"], i["batches"], i["samples"], 1
        self.table[self.rank].view(self.P, REC_WORDS).copy_(h, non_blocking=True)

    def set_root(self, li: int, root) -> None:
        if root is None:
            return
        if not isinstance(root, torch.Tensor):
            root = torch.frombuffer(bytearray(root), dtype=torch.uint8)
        self._rec(li)[R_ROOT:R_ROOT + 8].view(torch.uint8).copy_(root.reshape(-1), non_blocking=True)

    def set_f32(self, li: int, word: int, value) -> None:
        v = value if isinstance(value, torch.Tensor) else torch.tensor(float(value))
        self._rec(li)[word:word + 1].view(torch.float32).copy_(v.reshape(1).to(torch.float32), non_blocking=True)

    def set_i32(self, li: int, word: int, value) -> None:
        v = value if isinstance(value, torch.Tensor) else torch.tensor(int(value))
        self._rec(li)[word:word + 1].copy_(v.reshape(1).to(torch.int32), non_blocking=True)

    def exchange(self) -> None:
        self.comm.exchange_slots(self.table)

    def parse(self, host_table: torch.Tensor) -> List[Dict]:
        """``host_table``: the table as read back.  Returns one dict per valid peer, sorted by peer id."""
        t = host_table.reshape(self.world * self.P, REC_WORDS).contiguous()
        f = t.view(torch.float32)
        out = []
        for r in range(t.shape[0]):
            if int(t[r, R_VALID]) != 1:
                continue
            out.append({"peer": int(t[r, R_PEER]), "batches": int(t[r, R_BATCHES]), "samples": int(t[r, R_SAMPLES]),
                        "root": bytes(t[r, R_ROOT:R_ROOT + 8].contiguous().view(torch.uint8).numpy().tobytes()),
                        "loss": float(f[r, R_LOSS]), "local_correct": int(t[r, R_LCORRECT]), "local_total": int(t[r, R_LTOTAL]),
                        "eval_loss_sum": float(f[r, R_ELOSS]), "eval_correct": int(t[r, R_ECORRECT]),
                        "eval_n_ds": int(t[r, R_ENDS]), "eval_total": int(t[r, R_ETOTAL]),
                        "eval_batches": int(t[r, R_EBATCHES])})
        out.sort(key=lambda d: d["peer"])
        return out
