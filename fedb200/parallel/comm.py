"""Comm interface + loopback and torch.distributed implementations.

The reference has no communication layer at all: server mode pickles ndarrays through Ray's
object store, serverless mode appends to a Python list.  The logical collectives it
performs — broadcast global model, gather client weights, (weighted) mean, scalar metric reduce —
are the methods below.  ``LoopbackComm`` (world = 1) is the CPU plumbing backend; ``TorchDistComm``
is the NCCL (+ ATen) baseline the fused peer-memory kernels are measured against
(:class:`fedb200.parallel.peer_mem.PeerComm` implements the same interface with one-kernel
collectives over NVLink).
"""
from __future__ import annotations

import os
from typing import Any, List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from .placement import owner_of, peers_per_rank


def init_distributed(device: Optional[torch.device] = None) -> tuple:
    """Join the torchrun rendezvous if there is one. Returns (rank, world, local_rank)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        use_cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        if use_cuda:
            torch.cuda.set_device(local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        else:
            dist.init_process_group("gloo")
    return rank, world, local


class Comm:
    rank: int = 0
    world: int = 1
    name: str = "base"
    executes_relays: bool = False   # True: two-hop routes of a RelayPlan really move through a relay GPU (peer backend)
    launches: int = 0          # number of this repo's kernels launched by the comm layer
    placement = None           # peer -> (rank, slot) table set by the runtime (None: block placement)

    def owner(self, peer: int, num_peers: int):
        return self.placement.owner(peer) if self.placement is not None else owner_of(peer, num_peers, self.world)

    def __init__(self, device):
        self.device = torch.device(device)

    # -- memory -----------------------------------------------------------------------------
    def nvls_available(self) -> bool:
        return False

    def alloc(self, name: str, numel: int, dtype=torch.float32, multicast: bool = False) -> torch.Tensor:
        """Buffer every rank allocates identically (peer-mapped on the peer backend; ``multicast``: also bound to an NVLS
        multicast object there)."""
        return torch.zeros(numel, dtype=dtype, device=self.device)

    def delta_buffer(self, rows: int, D: int, dtype=torch.float32) -> torch.Tensor:
        """Zeroed ``[rows, D_pad]`` buffer for the update scorer; rows are padded to the segment layout of the Gram
        kernel (:func:`fedb200.ops.cuda_ops.gram_layout`), the caller writes ``[:, :D]``."""
        from ..ops.cuda_ops import gram_layout
        self._gram_layout = gram_layout(D, self.world, rows)
        Dp = self._gram_layout["D_pad"]
        return self.alloc("deltas", rows * Dp, dtype).view(rows, Dp)

    # -- control ----------------------------------------------------------------------------
    def barrier(self) -> None:
        pass

    def all_gather_object(self, obj: Any) -> List[Any]:
        return [obj]

    def reduce_scalars(self, values: Sequence[float], op: str = "sum") -> List[float]:
        return [float(v) for v in values]

    # -- collectives over flat arenas -----------------------------------------------------------
    def weighted_mean(self, acc: torch.Tensor, weight: float, out: torch.Tensor, shadow=None, **kw) -> None:
        """``out = sum_r acc_r / sum_r weight_r`` on every rank (``acc_r`` = local sum n_k w_k);
        ``shadow`` optionally receives the bf16 copy in the same pass."""
        raise NotImplementedError

    def exchange_slots(self, table: torch.Tensor) -> None:
        """All-gather of small per-rank records: ``table`` is ``[world, k]`` with row ``rank`` filled locally; afterwards
        every rank holds every row.  World 1: nothing to do."""
        return None

    def mix_rows(self, slots: torch.Tensor, w_rows: np.ndarray, out: torch.Tensor, num_peers: int, plan=None) -> None:
        """``out[i] = sum_j w_rows[i, j] * slot(j)`` for the local peers i; ``slots`` is the local
        ``[P_local, D]`` symmetric buffer, j ranges over all ``num_peers`` global peers."""
        raise NotImplementedError

    def gram(self, deltas: torch.Tensor, num_peers: int) -> torch.Tensor:
        """Global ``[N, N]`` Gram matrix of all peers' delta rows (fp32, on every rank)."""
        raise NotImplementedError

    # -- async executor hooks (single-process backends: stream order already provides the ordering) ----------
    def peer_ptr(self, t: torch.Tensor, rank: int) -> int:
        return t.data_ptr()

    def flag_set(self, flags: torch.Tensor, index: int, value: int) -> None:
        flags[index] = value

    def flag_wait(self, flags: torch.Tensor, owner_rank: int, index: int, value: int) -> None:
        pass

    def check_errors(self) -> None:
        pass

    def error_word(self) -> Optional[torch.Tensor]:
        """Device word that a timed-out spin of the peer protocol sets (None: the backend has no device-side waits).  The
        round loops read it with the round's other results — no extra synchronisation."""
        return None

    def close(self) -> None:
        pass


class LoopbackComm(Comm):
    """Single-process backend: every peer is local.  Uses the repo's CUDA kernels when the
    buffers live on a GPU, plain torch on the CPU."""
    name = "loopback"

    def weighted_mean(self, acc, weight, out, shadow=None, **kw):
        if float(weight) <= 0.0:              # every contributor masked out: keep the previous global model
            return
        if acc.is_cuda:
            from ..ops import cuda_ops
            cuda_ops.mix_n([acc], [1.0 / float(weight)], out, shadow)
            self.launches += 1
        else:
            torch.mul(acc, 1.0 / float(weight), out=out)
            if shadow is not None:
                shadow.copy_(out)

    def mix_rows(self, slots, w_rows, out, num_peers, plan=None):
        P = slots.shape[0]
        assert P == num_peers, "loopback holds every peer locally"
        slot = lambda j: slots[self.owner(j, num_peers)[1]]
        for i in range(P):
            js = [j for j in range(P) if w_rows[i, j] != 0.0]
            if slots.is_cuda:
                from ..ops import cuda_ops
                cuda_ops.mix_n([slot(j) for j in js], [float(w_rows[i, j]) for j in js], out[i])
                self.launches += 1
            else:
                out[i].zero_()
                for j in js:
                    out[i].add_(slot(j), alpha=float(w_rows[i, j]))

    def gram(self, deltas, num_peers):
        if deltas.is_cuda:
            from ..ops import cuda_ops
            self.launches += 1
            lay = getattr(self, "_gram_layout", None)
            if lay is not None and lay["S"] > 1 and deltas.shape[1] == lay["D_pad"]:
                out = torch.zeros(deltas.shape[0], deltas.shape[0], dtype=torch.float32, device=deltas.device)
                return cuda_ops.gram_seg([deltas.data_ptr()], deltas.shape[0], lay, 0, out)
            return cuda_ops.gram(deltas)
        d = deltas.to(torch.float32)
        return d @ d.T


class TorchDistComm(Comm):
    """NCCL / gloo baseline: library collectives + ATen elementwise ops."""
    name = "nccl"

    def __init__(self, device):
        super().__init__(device)
        assert dist.is_initialized()
        self.rank, self.world = dist.get_rank(), dist.get_world_size()

    def barrier(self):
        dist.barrier()

    def all_gather_object(self, obj):
        out = [None] * self.world
        dist.all_gather_object(out, obj)
        return out

    def reduce_scalars(self, values, op="sum"):
        t = torch.tensor(list(values), dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX)
        return t.tolist()

    def exchange_slots(self, table):
        rows = [torch.empty_like(table[0]) for _ in range(self.world)]
        dist.all_gather(rows, table[self.rank].contiguous())
        for r, row in enumerate(rows):
            table[r].copy_(row)

    def weighted_mean(self, acc, weight, out, shadow=None, **kw):
        w = torch.tensor([float(weight)], dtype=torch.float64, device=self.device)
        dist.all_reduce(w)
        if float(w.item()) <= 0.0:            # every contributor masked out: keep the previous global model
            return
        if out.data_ptr() != acc.data_ptr():
            out.copy_(acc)
        dist.all_reduce(out)
        out.mul_(1.0 / float(w.item()))
        if shadow is not None:
            shadow.copy_(out)

    def mix_rows(self, slots, w_rows, out, num_peers, plan=None):
        ppr = peers_per_rank(num_peers, self.world)
        gathered = [torch.empty_like(slots) for _ in range(self.world)]
        dist.all_gather(gathered, slots.contiguous())
        out.zero_()
        for i in range(out.shape[0]):
            for j in range(num_peers):
                w = float(w_rows[i, j])
                if w != 0.0:
                    r, l = self.owner(j, num_peers)
                    out[i].add_(gathered[r][l], alpha=w)

    def gram(self, deltas, num_peers):
        ppr = peers_per_rank(num_peers, self.world)
        d = deltas.to(torch.float32)
        # shard the reduction dimension: every rank needs all rows of its K-slice -> all_gather rows
        gathered = [torch.empty_like(d) for _ in range(self.world)]
        dist.all_gather(gathered, d.contiguous())
        allrows = torch.cat(gathered, dim=0)[:num_peers] if ppr * self.world == num_peers else \
            torch.cat([g for g in gathered], dim=0)[:num_peers]
        return allrows @ allrows.T


def make_comm(kind: str, device, rank: int = 0, world: int = 1) -> Comm:
    device = torch.device(device)
    if kind == "auto":
        kind = "loopback" if world == 1 else ("peer" if device.type == "cuda" else "nccl")
    if kind == "loopback":
        if world != 1:
            raise ValueError("loopback backend is single-process")
        return LoopbackComm(device)
    if kind == "nccl":
        return TorchDistComm(device) if world > 1 else LoopbackComm(device)
    if kind == "peer":
        if world == 1:
            return LoopbackComm(device)
        from .peer_mem import PeerComm
        return PeerComm(device)
    raise ValueError(kind)
