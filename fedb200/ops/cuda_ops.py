"""Python wrappers (torch tensors in, kernels out) over the C ABI in ``include/fedb200.h``.

Every wrapper validates dtype/device/contiguity, passes raw pointers plus the current CUDA stream, and
bumps :data:`fedb200.ops._lib.LAUNCHES` by the number of kernels it launches.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import call, count

EPI_BIAS, EPI_BIAS_DROP_RES, EPI_BIAS_GELU, EPI_MUL_GELU_GRAD, EPI_ADD, EPI_ATOMIC_F32 = range(6)
LAYOUT_NT, LAYOUT_NN, LAYOUT_TN = 0, 1, 2


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _chk(t: torch.Tensor, dtype, name: str) -> None:
    if not t.is_cuda or t.dtype != dtype:
        raise TypeError(f"{name}: expected cuda {dtype}, got {t.device} {t.dtype}")


# ---- flat arena -----------------------------------------------------------------------------------------
def adamw_flat(p, g, m, v, shadow, step, lr, b1, b2, eps, wd, grad_scale=1.0, zero_grad=True) -> None:
    for t, n in ((p, "p"), (g, "g"), (m, "m"), (v, "v")):
        _chk(t, torch.float32, n)
    if shadow is not None:
        _chk(shadow, torch.bfloat16, "shadow")
    call("fedb200_adamw_flat", _p(p), _p(g), _p(m), _p(v), _p(shadow), p.numel(), int(step), lr, b1, b2, eps, wd,
         grad_scale, int(zero_grad), _stream())
    count()


MAX_MIX_SOURCES = 64          # FEDB200_MAX_MIX_SOURCES in include/fedb200.h


def mix_n(srcs: Sequence, weights: Sequence[float], out: torch.Tensor, shadow: Optional[torch.Tensor] = None,
          numel: Optional[int] = None) -> None:
    """``out = sum_j weights[j] * srcs[j]``; a source may be a tensor or a raw (peer-mapped) device pointer.  More than
    ``FEDB200_MAX_MIX_SOURCES`` (64) sources are combined in passes: ``out`` re-enters the next pass as a source with weight 1
    (each thread reads and then writes its own elements, so the in-place accumulation is safe)."""
    n = len(srcs)
    if n > MAX_MIX_SOURCES:
        srcs, weights = list(srcs), list(weights)
        mix_n(srcs[:MAX_MIX_SOURCES], weights[:MAX_MIX_SOURCES], out, None, numel)
        o = MAX_MIX_SOURCES
        while o < n:
            e = min(n, o + MAX_MIX_SOURCES - 1)
            mix_n([out] + srcs[o:e], [1.0] + weights[o:e], out, shadow if e == n else None, numel)
            o = e
        return
    ptrs = (C.c_void_p * n)(*[(s.data_ptr() if isinstance(s, torch.Tensor) else int(s)) for s in srcs])
    ws = (C.c_float * n)(*[float(w) for w in weights])
    _chk(out, torch.float32, "out")
    call("fedb200_mix_n", C.cast(ptrs, C.c_void_p), C.cast(ws, C.c_void_p), n, _p(out), _p(shadow),
         numel if numel is not None else out.numel(), _stream())
    count()


def _ptr(x) -> int:
    return x.data_ptr() if isinstance(x, torch.Tensor) else int(x)


def gossip_mix_async(self_src: torch.Tensor, out: torch.Tensor, rings: Sequence, chunk_flags: Sequence, pubs: Sequence,
                     src_peers: Sequence[int], mask: Optional[torch.Tensor], decision: torch.Tensor, log: torch.Tensor,
                     err, R: int, chunk_elems: int, my_epoch: int, launch_id: int, stale_w: Sequence[float],
                     max_ctas: int = 0, timeout_ms: int = 20000) -> None:
    """One asynchronous mix from observed versions (``csrc/gossip.cu::gossip_mix_async_kernel``); pointers may be peer-mapped."""
    n = len(rings)
    arr = lambda xs: C.cast((C.c_void_p * max(n, 1))(*[_ptr(x) for x in xs]), C.c_void_p)
    peers = (C.c_int * max(n, 1))(*[int(p) for p in src_peers])
    sw = (C.c_float * 32)(*[float(w) for w in stale_w])
    call("fedb200_gossip_mix_async", _p(self_src), _p(out), arr(rings), arr(chunk_flags), arr(pubs), C.cast(peers, C.c_void_p),
         n, _p(mask), _p(decision), _p(log), _ptr(err), R, out.numel(), chunk_elems, int(my_epoch), int(launch_id),
         C.cast(sw, C.c_void_p), max_ctas, timeout_ms, _stream())
    count()


def copy_flagged(src: torch.Tensor, dst, chunk_elems: int, flags, value: int) -> None:
    call("fedb200_copy_flagged", _p(src), _ptr(dst), src.numel(), chunk_elems, _ptr(flags), int(value), _stream())
    count()


def set_flags(flags, value: int, n: int) -> None:
    call("fedb200_set_flags", _ptr(flags), int(value), n, _stream())
    count()


def colsum_bf16(x: torch.Tensor, out: torch.Tensor) -> None:
    _chk(x, torch.bfloat16, "x")
    _chk(out, torch.float32, "out")
    M, N = x.shape
    call("fedb200_colsum_bf16", _p(x), _p(out), M, N, x.stride(0), _stream())
    count()


def cast_bf16(src: torch.Tensor, dst: torch.Tensor) -> None:
    _chk(src, torch.float32, "src")
    _chk(dst, torch.bfloat16, "dst")
    call("fedb200_cast_f32_bf16", _p(src), _p(dst), src.numel(), _stream())
    count()


def delta_bf16(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor) -> None:
    call("fedb200_delta_bf16", _p(a), _p(b), _p(out), a.numel(), _stream())
    count()


# ---- trainer ops -------------------------------------------------------------------------------------------
def gemm(layout: int, A: torch.Tensor, B: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int, epi: int = EPI_BIAS,
         bias: Optional[torch.Tensor] = None, aux: Optional[torch.Tensor] = None, out2: Optional[torch.Tensor] = None,
         act: int = 0, seed: int = 0, dropout_p: float = 0.0, splits: int = 1, max_ctas: int = 0) -> None:
    _chk(A, torch.bfloat16, "A")
    _chk(B, torch.bfloat16, "B")
    call("fedb200_gemm_bf16", layout, _p(A), A.stride(0), _p(B), B.stride(0), M, N, K, epi, _p(out), _p(out2),
         out.stride(0), _p(bias), _p(aux), aux.stride(0) if aux is not None else 0, act, seed, dropout_p, splits,
         max_ctas, _stream())
    count()


def layernorm_fwd(z, gamma, beta, y, mean, rstd, eps=1e-12) -> None:
    T, H = z.shape
    call("fedb200_layernorm_fwd", _p(z), _p(gamma), _p(beta), _p(y), _p(mean), _p(rstd), T, H, eps, _stream())
    count()


def layernorm_bwd(dy, z, mean, rstd, gamma, dz, df, dgamma, dbeta, dbias, seed=0, dropout_p=0.0) -> None:
    T, H = z.shape
    call("fedb200_layernorm_bwd", _p(dy), _p(z), _p(mean), _p(rstd), _p(gamma), _p(dz), _p(df), _p(dgamma), _p(dbeta),
         _p(dbias), T, H, seed, dropout_p, _stream())
    count()


def embed_fwd(ids, pos, tt, word, posw, typew, gamma, beta, out, mean, rstd, eps=1e-12, seed=0, dropout_p=0.0) -> None:
    T, H = out.shape
    call("fedb200_embed_fwd", _p(ids), _p(pos), _p(tt), _p(word), _p(posw), _p(typew), _p(gamma), _p(beta), _p(out),
         _p(mean), _p(rstd), T, H, eps, seed, dropout_p, _stream())
    count()


def embed_bwd(dx, ids, pos, tt, word, posw, typew, gamma, mean, rstd, gword, gpos, gtype, dgamma, dbeta, seed=0,
              dropout_p=0.0) -> None:
    T, H = dx.shape
    call("fedb200_embed_bwd", _p(dx), _p(ids), _p(pos), _p(tt), _p(word), _p(posw), _p(typew), _p(gamma), _p(mean),
         _p(rstd), _p(gword), _p(gpos), _p(gtype), _p(dgamma), _p(dbeta), T, H, typew.shape[0], seed, dropout_p,
         _stream())
    count()


def head_fwd(x, first_tok, Wp, bp, Wc, bc, labels, pooled, pooled_drop, logits, loss, pred, dlogits, grad_scale,
             seed=0, dropout_p=0.0) -> None:
    B, C_ = logits.shape
    H = x.shape[1]
    call("fedb200_head_fwd", _p(x), _p(first_tok), _p(Wp), _p(bp), _p(Wc), _p(bc), _p(labels), _p(pooled),
         _p(pooled_drop), _p(logits), _p(loss), _p(pred), _p(dlogits), B, H, C_, grad_scale, seed, dropout_p, _stream())
    count(2)


def head_bwd(x, first_tok, Wp, Wc, pooled, pooled_drop, dlogits, dpre, gWp, gbp, gWc, gbc, dx, seed=0,
             dropout_p=0.0) -> None:
    B, C_ = dlogits.shape
    H = x.shape[1]
    call("fedb200_head_bwd", _p(x), _p(first_tok), _p(Wp), _p(Wc), _p(pooled), _p(pooled_drop), _p(dlogits), _p(dpre),
         _p(gWp), _p(gbp), _p(gWc), _p(gbc), _p(dx), B, H, C_, seed, dropout_p, _stream())
    count(3)


def attn_fwd(qkv, ctx, lse, cu_seqlens, B, heads, max_seqlen, scale, seed=0, dropout_p=0.0, drop_bits=None) -> None:
    """``drop_bits`` (int32/uint32 ``[heads, T, ceil(max_seqlen/32)]``, optional) receives the dropout keep-bits."""
    T = qkv.shape[0]
    call("fedb200_attn_fwd", _p(qkv), _p(ctx), _p(lse), _p(cu_seqlens), B, T, heads, max_seqlen, scale, seed,
         dropout_p, _p(drop_bits), _stream())
    count()


def attn_bwd(qkv, ctx, dctx, lse, delta, dqkv, dq_accum, cu_seqlens, B, heads, max_seqlen, scale, seed=0,
             dropout_p=0.0, dbias=None, drop_bits=None) -> None:
    """``dq_accum`` is fp32 scratch of ``ceil(max_seqlen / 128) * T * H`` floats (per-kv-tile dQ partials, no
    initialisation needed); ``dbias`` (fp32 [3H], optional) receives the q/k/v bias gradients from the fused
    finishing pass."""
    T = qkv.shape[0]
    call("fedb200_attn_bwd", _p(qkv), _p(ctx), _p(dctx), _p(lse), _p(delta), _p(dqkv), _p(dq_accum), _p(dbias),
         _p(cu_seqlens), B, T, heads, max_seqlen, scale, seed, dropout_p, _p(drop_bits), _stream())
    count(3)


# ---- federated ops --------------------------------------------------------------------------------------------
def gram(deltas: torch.Tensor, col_begin: int = 0, col_end: Optional[int] = None, out: Optional[torch.Tensor] = None,
         peer_ptrs: Optional[Sequence[int]] = None, rows_per_rank: Optional[int] = None, max_ctas: int = 0
         ) -> torch.Tensor:
    """Gram matrix of bf16 rows.  Single buffer by default; ``peer_ptrs`` lists every rank's (peer-mapped)
    ``[rows_per_rank, D]`` buffer for the multi-GPU fused form."""
    _chk(deltas, torch.bfloat16, "deltas")
    rows, D = deltas.shape
    ptr_list = [deltas.data_ptr()] if peer_ptrs is None else [int(p) for p in peer_ptrs]
    rpr = rows if rows_per_rank is None else rows_per_rank
    n = len(ptr_list) * rpr
    if out is None:
        out = torch.zeros(n, n, dtype=torch.float32, device=deltas.device)
    ptrs = (C.c_void_p * len(ptr_list))(*ptr_list)
    call("fedb200_gram_bf16", C.cast(ptrs, C.c_void_p), len(ptr_list), rpr, D, deltas.stride(0), col_begin,
         D if col_end is None else col_end, _p(out), max_ctas, _stream())
    count()
    return out


def gram_layout(D: int, n_ranks: int, rows_per_rank: int) -> dict:
    """Padded row layout for the segment-stacked Gram kernel: every client row is ``nseg_total`` segments of
    ``seg_len`` elements (zero padded to ``D_pad``); rank r reduces segments ``[r*S, (r+1)*S)``."""
    clients = max(1, n_ranks * rows_per_rank)
    S = 1
    while S * 2 * clients <= 128:
        S *= 2
    seg_len = (D + n_ranks * S * 64 - 1) // (n_ranks * S * 64) * 64
    return {"S": S, "seg_len": seg_len, "nseg_total": n_ranks * S, "D_pad": n_ranks * S * seg_len}


def gram_seg(ptrs: Sequence[int], rows_per_rank: int, layout: dict, seg_begin: int, out: torch.Tensor,
             max_ctas: int = 0) -> torch.Tensor:
    """Segment-stacked Gram over (peer) buffers laid out by :func:`gram_layout`; accumulates into ``out``."""
    arr = (C.c_void_p * len(ptrs))(*[int(p) for p in ptrs])
    call("fedb200_gram_bf16_seg", C.cast(arr, C.c_void_p), len(ptrs), rows_per_rank, layout["seg_len"], layout["nseg_total"],
         seg_begin, layout["S"], _p(out), max_ctas, _stream())
    count()
    return out


def pagerank(mat: torch.Tensor, mode: int = 0, alpha: float = 0.85, tol: float = 1e-6, max_iter: int = 100,
             k_sigma: float = 1.0, low_only: bool = False) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    _chk(mat, torch.float32, "mat")
    n = mat.shape[0]
    ranks = torch.empty(n, dtype=torch.float64, device=mat.device)
    mask = torch.empty(n, dtype=torch.int32, device=mat.device)
    info = torch.empty(4, dtype=torch.float64, device=mat.device)
    call("fedb200_pagerank", _p(mat.contiguous()), n, mode, alpha, tol, max_iter, k_sigma, int(low_only), _p(ranks),
         _p(mask), _p(info), _stream())
    count()
    return ranks, mask, info


def sha256_leaves(src: torch.Tensor, dst: Optional[torch.Tensor], chunk_bytes: int, leaves: Optional[torch.Tensor] = None,
                  dst_ptr: Optional[int] = None) -> torch.Tensor:
    nbytes = src.numel() * src.element_size()
    n_chunks = (nbytes + chunk_bytes - 1) // chunk_bytes
    if leaves is None:
        leaves = torch.empty(n_chunks, 32, dtype=torch.uint8, device=src.device)
    d = dst_ptr if dst_ptr is not None else _p(dst)
    call("fedb200_sha256_leaves", _p(src), d, nbytes, chunk_bytes, _p(leaves), _stream())
    count()
    return leaves


def merkle_root(leaves: torch.Tensor, scratch: Optional[torch.Tensor] = None, root: Optional[torch.Tensor] = None
                ) -> torch.Tensor:
    """Reduces (and clobbers) ``leaves`` [n, 32] uint8; returns the 32-byte root as a device tensor."""
    n = leaves.shape[0]
    if scratch is None:
        scratch = torch.empty((n + 1) // 2 + 1, 32, dtype=torch.uint8, device=leaves.device)
    if root is None:
        root = torch.empty(32, dtype=torch.uint8, device=leaves.device)
    call("fedb200_merkle_root", _p(leaves), _p(scratch), n, _p(root), _stream())
    count(max(1, int(np.ceil(np.log2(max(n, 2))))))
    return root


_SHA_WS = {}


def sha256_root_static(src, nbytes: int, chunk_bytes: int, root_out: torch.Tensor) -> None:
    """Merkle root of ``nbytes`` at ``src`` (tensor or raw, possibly peer-mapped, device pointer) into ``root_out`` (32-byte
    device tensor) with NO workspace: the library's static device array holds the leaves and levels.  This is the call the
    receive path uses to check an update it is about to mix against the root its sender recorded — no allocation there."""
    call("fedb200_sha256_root_static", _ptr(src), int(nbytes), int(chunk_bytes), _p(root_out), _stream())
    count(2)


def sha256_copy_root(src: torch.Tensor, dst: Optional[torch.Tensor], chunk_bytes: int = 4096, to_host: bool = True):
    """Copy ``src`` -> ``dst`` hashing on the way; returns the Merkle root as host bytes (syncs the stream) or, with
    ``to_host=False``, as a fresh 32-byte device tensor (no sync)."""
    nbytes = src.numel() * src.element_size()
    n_chunks = (nbytes + chunk_bytes - 1) // chunk_bytes
    key = (src.device, n_chunks)
    if key not in _SHA_WS:
        _SHA_WS[key] = (torch.empty(n_chunks, 32, dtype=torch.uint8, device=src.device),
                        torch.empty((n_chunks + 1) // 2 + 1, 32, dtype=torch.uint8, device=src.device),
                        torch.empty(32, dtype=torch.uint8, device=src.device))
    leaves, scratch, root = _SHA_WS[key]
    sha256_leaves(src, dst, chunk_bytes, leaves)
    merkle_root(leaves, scratch, root)
    if not to_host:
        return root.clone()
    return bytes(root.cpu().numpy().tobytes())


def block_validate(headers: torch.Tensor, recorded: Optional[torch.Tensor] = None) -> Tuple[int, torch.Tensor]:
    """``headers``: uint8 [n, 128] on device.  Returns (first bad index or -1, hashes [n, 32])."""
    n = headers.shape[0]
    hashes = torch.empty(n, 32, dtype=torch.uint8, device=headers.device)
    bad = torch.empty(1, dtype=torch.int32, device=headers.device)
    call("fedb200_block_validate", _p(headers.contiguous()), _p(recorded), n, _p(hashes), _p(bad), _stream())
    count(2)
    b = int(bad.item())
    return (-1 if b < 0 or b >= n else b), hashes
