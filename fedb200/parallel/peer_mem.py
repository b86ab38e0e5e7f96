"""NVLink peer-memory communication backend: CUDA-IPC symmetric buffers + one-kernel
collectives.  One process per GPU; ``torch.distributed`` (NCCL) is used only to bootstrap — exchange the
64-byte IPC handles and small Python objects — every byte of model data moves through the hand-written
kernels in ``csrc/peer.cu`` / ``csrc/elementwise.cu`` / ``csrc/gram_tcgen05.cu`` reading and writing peer
memory directly.  The reference has no counterpart (no distributed communication at all)."""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops._lib import call, count
from .comm import Comm
from .placement import owner_of, peers_per_rank

FLAG_WORDS = 128              # FEDB200_FLAG_WORDS (include/fedb200.h)


class _RawCuda:
    """Minimal ``__cuda_array_interface__`` carrier so a raw device pointer becomes a torch tensor."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class SymmetricBuffer:
    def __init__(self, name: str, tensor: torch.Tensor, ptrs: List[int], nbytes: int):
        self.name, self.tensor, self.ptrs, self.nbytes = name, tensor, ptrs, nbytes

    def table(self, byte_offset: int = 0):
        arr = (C.c_void_p * len(self.ptrs))(*[p + byte_offset for p in self.ptrs])
        return arr


class PeerComm(Comm):
    name = "peer"

    def __init__(self, device, timeout_ms: int = 20000):
        super().__init__(device)
        assert dist.is_initialized(), "PeerComm needs torch.distributed for the handle exchange"
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        if self.world > 8:
            raise ValueError("the peer backend addresses one NVSwitch box (<= 8 GPUs)")
        self.timeout_ms = timeout_ms
        self.bufs: Dict[str, SymmetricBuffer] = {}
        self._by_ptr: Dict[int, Tuple[SymmetricBuffer, int]] = {}
        self._own: List[int] = []
        self._mapped: List[int] = []
        self.epoch = 0
        self.os_epoch = 0
        self.xchg_epoch = 0
        self.bar_epoch = 0
        self.flags = self._alloc_symmetric("__flags__", FLAG_WORDS * 4)
        self.gram_part = None

    # ---- symmetric memory -------------------------------------------------------------------------------
    def _alloc_symmetric(self, name: str, nbytes: int) -> SymmetricBuffer:
        nbytes = (nbytes + 255) // 256 * 256
        ptr = C.c_void_p()
        handle = C.create_string_buffer(64)
        call("fedb200_peer_alloc", nbytes, C.byref(ptr), handle)
        self._own.append(ptr.value)
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw))
        ptrs = []
        for r, h in enumerate(handles):
            if r == self.rank:
                ptrs.append(ptr.value)
            else:
                mp = C.c_void_p()
                call("fedb200_peer_open", C.create_string_buffer(h, 64), C.byref(mp))
                self._mapped.append(mp.value)
                ptrs.append(mp.value)
        t = torch.as_tensor(_RawCuda(ptr.value, nbytes), device=self.device)
        buf = SymmetricBuffer(name, t, ptrs, nbytes)
        while name in self.bufs:                  # a second runtime on the same backend re-uses buffer names
            name += "'"
        self.bufs[name] = buf
        dist.barrier()
        return buf

    def alloc(self, name: str, numel: int, dtype=torch.float32) -> torch.Tensor:
        esz = torch.empty(0, dtype=dtype).element_size()
        buf = self._alloc_symmetric(name, numel * esz)
        view = buf.tensor[: numel * esz].view(dtype)
        self._by_ptr[view.data_ptr()] = (buf, 0)
        return view

    def _locate(self, t: torch.Tensor) -> Tuple[SymmetricBuffer, int]:
        """(symmetric buffer, byte offset) containing tensor ``t``."""
        p = t.data_ptr()
        for buf in self.bufs.values():
            base = buf.ptrs[self.rank]
            if base <= p < base + buf.nbytes:
                return buf, p - base
        raise ValueError("tensor does not live in a symmetric (peer-mapped) buffer; allocate it with comm.alloc()")

    # ---- control ----------------------------------------------------------------------------------------------
    def barrier(self) -> None:
        dist.barrier()

    def device_barrier(self) -> None:
        """Cross-rank barrier as a 1-CTA kernel on the current stream (no host involvement)."""
        self.bar_epoch += 1
        call("fedb200_peer_barrier", C.cast(self.flags.table(), C.c_void_p), self.rank, self.world, self.bar_epoch,
             self.timeout_ms, torch.cuda.current_stream().cuda_stream)
        count()
        self.launches += 1

    def all_gather_object(self, obj):
        out = [None] * self.world
        dist.all_gather_object(out, obj)
        return out

    def reduce_scalars(self, values, op="sum"):
        t = torch.tensor(list(values), dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX)
        return t.tolist()

    def peer_ptr(self, t: torch.Tensor, rank: int) -> int:
        """Address of the same symmetric location as ``t`` inside ``rank``'s copy of the buffer."""
        buf, off = self._locate(t)
        return buf.ptrs[rank] + off

    def flag_set(self, flags: torch.Tensor, index: int, value: int) -> None:
        call("fedb200_flag_set", flags.data_ptr() + 4 * index, int(value), torch.cuda.current_stream().cuda_stream)
        count()
        self.launches += 1

    def flag_wait(self, flags: torch.Tensor, owner_rank: int, index: int, value: int) -> None:
        """Block the current stream until ``owner_rank``'s copy of ``flags[index]`` has reached ``value``."""
        err = self.flags.tensor.data_ptr() + 4 * 18
        call("fedb200_flag_wait", self.peer_ptr(flags, owner_rank) + 4 * index, int(value), self.timeout_ms, err,
             torch.cuda.current_stream().cuda_stream)
        count()
        self.launches += 1

    def error_word(self) -> torch.Tensor:
        return self.flags.tensor.view(torch.int32)[18:19]

    def check_errors(self) -> None:
        if int(self.flags.tensor.view(torch.int32)[18].item()) != 0:
            raise RuntimeError("a peer-memory barrier timed out (flag protocol error or a rank is missing)")

    # ---- collectives ---------------------------------------------------------------------------------------------
    def weighted_mean(self, acc: torch.Tensor, weight: float, out: torch.Tensor, shadow: Optional[torch.Tensor] = None,
                      inv_total: float = 0.0, oneshot: Optional[bool] = None, max_ctas: int = 0) -> None:
        """Fused FedAvg over peer memory.  ``oneshot`` (default: payloads <= 64 KB) reads every peer's whole buffer and
        writes only the local ``out``; otherwise two-shot reduce-scatter + push.  ``max_ctas`` caps the grid for overlap
        with compute on a side stream."""
        ab, ao = self._locate(acc)
        ob, oo = self._locate(out)
        sh = None
        if shadow is not None:
            sb, so = self._locate(shadow)
            sh = C.cast(sb.table(so), C.c_void_p)
        if oneshot is None:
            oneshot = acc.numel() * 4 <= (64 << 10)
        if oneshot:
            self.os_epoch += 1
            ep = self.os_epoch
        else:
            self.epoch += 1
            ep = self.epoch
        call("fedb200_fedavg_reduce_ex", C.cast(ab.table(ao), C.c_void_p), C.cast(ob.table(oo), C.c_void_p), sh,
             C.cast(self.flags.table(), C.c_void_p), self.rank, self.world, acc.numel(), float(weight), float(inv_total),
             ep, self.timeout_ms, int(bool(oneshot)), int(max_ctas), torch.cuda.current_stream().cuda_stream)
        count()
        self.launches += 1

    def exchange_slots(self, table: torch.Tensor) -> None:
        """``table`` is a symmetric ``[world, k]`` buffer (``comm.alloc``) whose row ``rank`` this rank has filled; after the
        call (stream-ordered, no host involvement) every rank's copy holds every rank's row."""
        tb, to = self._locate(table)
        slot_bytes = table[0].numel() * table.element_size()
        self.xchg_epoch += 1
        call("fedb200_peer_exchange", C.cast(tb.table(to), C.c_void_p), C.cast(self.flags.table(), C.c_void_p), self.rank,
             self.world, slot_bytes, self.xchg_epoch, self.timeout_ms, torch.cuda.current_stream().cuda_stream)
        count()
        self.launches += 1

    def mix_rows(self, slots: torch.Tensor, w_rows: np.ndarray, out: torch.Tensor, num_peers: int) -> None:
        from ..ops import cuda_ops
        sb, so = self._locate(slots)
        D = slots.shape[1]
        self.device_barrier()                                   # every peer's slot is published
        for i in range(out.shape[0]):
            srcs, ws = [], []
            for j in range(num_peers):
                w = float(w_rows[i, j])
                if w != 0.0:
                    r, l = self.owner(j, num_peers)
                    srcs.append(sb.ptrs[r] + so + l * D * 4)
                    ws.append(w)
            if srcs:
                cuda_ops.mix_n(srcs, ws, out[i])
                self.launches += 1
        self.device_barrier()                                   # nobody overwrites a slot that is still being read

    def gram(self, deltas: torch.Tensor, num_peers: int) -> torch.Tensor:
        """All-gather-fused Gram: every rank reduces ITS column slice of every rank's rows (TMA reads over
        NVLink), then the tiny [N, N] partials are summed with the fused reduce."""
        db, do = self._locate(deltas)
        P, D = deltas.shape
        n = self.world * P
        if self.gram_part is None or self.gram_part.numel() != n * n:
            self.gram_part = self.alloc(f"gram_part_{n}", n * n)
            self.gram_out = self.alloc(f"gram_out_{n}", n * n)
        self.gram_part.zero_()
        self.device_barrier()                                   # all delta rows are written
        lay = getattr(self, "_gram_layout", None)
        if lay is not None and lay["S"] > 1 and D == lay["D_pad"]:
            from ..ops import cuda_ops
            cuda_ops.gram_seg([p + do for p in db.ptrs], P, lay, self.rank * lay["S"], self.gram_part)
        else:
            kb = (D + 63) // 64
            per = (kb + self.world - 1) // self.world
            lo, hi = min(D, self.rank * per * 64), min(D, (self.rank + 1) * per * 64)
            ptrs = (C.c_void_p * self.world)(*[p + do for p in db.ptrs])
            call("fedb200_gram_bf16", C.cast(ptrs, C.c_void_p), self.world, P, D, D, lo, hi, self.gram_part.data_ptr(), 0,
                 torch.cuda.current_stream().cuda_stream)
            count()
        self.launches += 1
        self.weighted_mean(self.gram_part, 1.0, self.gram_out, inv_total=1.0)
        return self.gram_out.view(n, n)

    def close(self) -> None:
        try:
            torch.cuda.synchronize()
            for p in self._mapped:
                _lib.lib().fedb200_peer_close(C.c_void_p(p))
            self._mapped.clear()
            if dist.is_initialized():
                dist.barrier()
            for p in self._own:
                _lib.lib().fedb200_peer_free(C.c_void_p(p))
            self._own.clear()
        except Exception:
            pass
