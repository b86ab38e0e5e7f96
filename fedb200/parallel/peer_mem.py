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
    def __init__(self, name: str, tensor: torch.Tensor, ptrs: List[int], nbytes: int, mc_ptr: int = 0, keep=None):
        self.name, self.tensor, self.ptrs, self.nbytes = name, tensor, ptrs, nbytes
        self.mc_ptr = mc_ptr            # multicast (NVLS) address of the same buffer, 0 if it is not bound to a multicast object
        self._keep = keep               # allocator handles that must outlive the buffer

    def table(self, byte_offset: int = 0):
        arr = (C.c_void_p * len(self.ptrs))(*[p + byte_offset for p in self.ptrs])
        return arr


class PeerComm(Comm):
    name = "peer"
    executes_relays = True

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
        self.relay_slots = 4                     # first-hop relay copies a rank can hold per gossip round (433 MB each)
        self.gram_epoch = 0                      # monotone over the life of the flag words, whatever N is

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

    def _alloc_multicast(self, name: str, nbytes: int) -> SymmetricBuffer:
        """Symmetric buffer bound to an NVLS multicast object.  Allocation, handle exchange and the multicast binding
        (CUDA VMM: cuMemCreate / cuMulticastCreate / AddDevice / BindMem) are host-side plumbing done by torch's
        symmetric-memory allocator; what comes back — every rank's unicast pointer and ONE multicast pointer — is what the
        kernels of this library consume.  Raises if the box exposes no multicast."""
        import torch.distributed._symmetric_memory as symm
        nbytes = (nbytes + 255) // 256 * 256
        t = symm.empty(nbytes, dtype=torch.uint8, device=self.device)
        t.zero_()
        h = symm.rendezvous(t, dist.group.WORLD)
        if int(h.multicast_ptr) == 0:
            raise RuntimeError("this box does not expose NVLS multicast (multicast_ptr == 0)")
        buf = SymmetricBuffer(name, t, [int(p) for p in h.buffer_ptrs], nbytes, int(h.multicast_ptr), keep=h)
        while name in self.bufs:
            name += "'"
        self.bufs[name] = buf
        torch.cuda.synchronize()
        dist.barrier()
        return buf

    def nvls_available(self) -> bool:
        if getattr(self, "_nvls_ok", None) is None:
            try:
                self._alloc_multicast("__nvls_probe__", 4096)
                self._nvls_ok = True
            except Exception:
                self._nvls_ok = False
        return self._nvls_ok

    def alloc(self, name: str, numel: int, dtype=torch.float32, multicast: bool = False) -> torch.Tensor:
        esz = torch.empty(0, dtype=dtype).element_size()
        buf = self._alloc_multicast(name, numel * esz) if multicast else self._alloc_symmetric(name, numel * esz)
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
                      inv_total: float = 0.0, oneshot: Optional[bool] = None, max_ctas: int = 0, nvls: bool = False) -> None:
        """Fused FedAvg over peer memory.  ``oneshot`` (default: payloads <= 64 KB) reads every peer's whole buffer and
        writes only the local ``out``; otherwise two-shot reduce-scatter + push.  ``max_ctas`` caps the grid for overlap
        with compute on a side stream.  ``nvls``: both buffers were allocated with ``multicast=True`` — the reduction is
        done inside the NVSwitch (``multimem.ld_reduce`` / ``multimem.st``), no bf16 shadow in that form."""
        ab, ao = self._locate(acc)
        ob, oo = self._locate(out)
        if nvls:
            if not (ab.mc_ptr and ob.mc_ptr):
                raise ValueError("nvls=True needs buffers allocated with comm.alloc(..., multicast=True)")
            if shadow is not None:
                raise ValueError("the NVLS form writes no bf16 shadow")
            self.epoch += 1
            call("fedb200_fedavg_reduce_nvls", C.cast(ab.table(ao), C.c_void_p), C.cast(ob.table(oo), C.c_void_p),
                 ab.mc_ptr + ao, ob.mc_ptr + oo, C.cast(self.flags.table(), C.c_void_p), self.rank, self.world, acc.numel(),
                 float(weight), float(inv_total), self.epoch, self.timeout_ms, int(max_ctas),
                 torch.cuda.current_stream().cuda_stream)
            count()
            self.launches += 1
            return
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

    def mix_rows(self, slots: torch.Tensor, w_rows: np.ndarray, out: torch.Tensor, num_peers: int, plan=None) -> None:
        """``out[i] = sum_j w_rows[i, j] * slot(j)`` for this rank's rows, ONE fused launch (``csrc/gossip.cu``): start
        barrier, optional first-hop relay copies + mid barrier, every distinct source read once for all local rows, end
        barrier.  ``plan`` (:class:`fedb200.p2p.gossip.RelayPlan`, identical on every rank) names the two-hop routes to
        execute through relay GPUs; ``row i`` of ``w_rows`` / ``out`` belongs to the rank's i-th local peer."""
        sb, so = self._locate(slots)
        D = slots.shape[1]
        rows = [i for i in range(out.shape[0]) if i < w_rows.shape[0] and np.any(w_rows[i] != 0.0)]
        local_ids = self.placement.local_ids(self.rank) if self.placement is not None else \
            [self.rank * slots.shape[0] + i for i in range(slots.shape[0])]
        relay_srcs, relay_dsts, any_relay = [], [], 0
        if plan is not None and plan.any:
            any_relay = 1
            if getattr(self, "relay_buf", None) is None or self.relay_buf.shape[1] != D:
                self.relay_buf = self.alloc("relay", self.relay_slots * D).view(self.relay_slots, D)
            for j, slot in plan.copies.get(self.rank, []):
                r, l = self.owner(j, num_peers)
                relay_srcs.append(sb.ptrs[r] + so + l * D * 4)
                relay_dsts.append(self.relay_buf[slot].data_ptr())
            rb, ro = self._locate(self.relay_buf)
        src_index, srcs = {}, []
        W = np.zeros((max(len(rows), 1), 0), dtype=np.float32)
        cols = []
        for ri, i in enumerate(rows):
            dst_peer = local_ids[i] if i < len(local_ids) else -1
            for j in range(num_peers):
                w = float(w_rows[i, j])
                if w == 0.0:
                    continue
                via = plan.via.get((dst_peer, j)) if plan is not None else None
                if via is not None:
                    key = ("relay",) + via
                    ptr = rb.ptrs[via[0]] + ro + via[1] * D * 4
                else:
                    r, l = self.owner(j, num_peers)
                    key = ("slot", r, l)
                    ptr = sb.ptrs[r] + so + l * D * 4
                if key not in src_index:
                    src_index[key] = len(srcs)
                    srcs.append(ptr)
                    cols.append(np.zeros(len(rows), dtype=np.float32))
                cols[src_index[key]][ri] += w
        if not rows:                       # nothing to mix on this rank: still take part in the barriers
            rows, srcs, cols = [0], [slots[0].data_ptr()], [np.zeros(1, dtype=np.float32)]
            keep = out[0].clone()
        if len(srcs) > 48:
            raise ValueError(f"{len(srcs)} distinct gossip sources on one rank exceed the fused kernel's 48; lower the topology degree")
        wmat = np.ascontiguousarray(np.stack(cols, axis=1), dtype=np.float32)      # [n_rows, n_src]
        self.mix_epoch = getattr(self, "mix_epoch", 0) + 1
        vp = lambda xs: C.cast((C.c_void_p * max(len(xs), 1))(*xs), C.c_void_p)
        call("fedb200_gossip_mix", vp(srcs), wmat.ctypes.data_as(C.c_void_p), len(srcs), vp([out[i].data_ptr() for i in rows]),
             len(rows), vp(relay_srcs), vp(relay_dsts), len(relay_srcs), any_relay, C.cast(self.flags.table(), C.c_void_p),
             self.rank, self.world, D, self.mix_epoch, self.timeout_ms, torch.cuda.current_stream().cuda_stream)
        count()
        self.launches += 1
        if not np.any(w_rows[: out.shape[0]] != 0.0):
            out[0].copy_(keep)

    def gram(self, deltas: torch.Tensor, num_peers: int) -> torch.Tensor:
        """All-gather-fused Gram in ONE launch: start barrier, TMA reads of this rank's column slice of every rank's rows
        over NVLink, tcgen05 accumulation, and the one-shot cross-rank sum of the tiny [N, N] partials all happen inside
        ``gram_tcgen05_kernel`` (round 1 needed zero_ + barrier kernel + Gram + two-shot reduce)."""
        db, do = self._locate(deltas)
        P, D = deltas.shape
        n = self.world * P
        if self.gram_part is None or self.gram_out.numel() != n * n:
            self.gram_part = self.alloc(f"gram_part_{n}", 2 * n * n)          # [2][n*n], zero-initialised by the allocator
            self.gram_out = self.alloc(f"gram_out_{n}", n * n)
        self.gram_epoch += 1
        pb, po = self._locate(self.gram_part)
        rows = (C.c_void_p * self.world)(*[p + do for p in db.ptrs])
        stream = torch.cuda.current_stream().cuda_stream
        lay = getattr(self, "_gram_layout", None)
        if lay is not None and lay["S"] > 1 and D == lay["D_pad"]:
            call("fedb200_gram_fused_seg", C.cast(rows, C.c_void_p), self.world, P, lay["seg_len"], lay["nseg_total"],
                 self.rank * lay["S"], lay["S"], C.cast(pb.table(po), C.c_void_p), C.cast(self.flags.table(), C.c_void_p),
                 self.gram_out.data_ptr(), self.rank, self.gram_epoch, self.timeout_ms, 0, stream)
        else:
            kb = (D + 63) // 64
            per = (kb + self.world - 1) // self.world
            lo, hi = min(D, self.rank * per * 64), min(D, (self.rank + 1) * per * 64)
            call("fedb200_gram_fused", C.cast(rows, C.c_void_p), self.world, P, D, D, lo, hi,
                 C.cast(pb.table(po), C.c_void_p), C.cast(self.flags.table(), C.c_void_p), self.gram_out.data_ptr(),
                 self.rank, self.gram_epoch, self.timeout_ms, 0, stream)
        count()
        self.launches += 1
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
            self.bufs.clear()                      # multicast buffers: tensors + handles released with their references
        except Exception:
            pass
