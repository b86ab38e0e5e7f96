"""ctypes loader for ``fedb200/lib/libfedb200.so`` (built in-tree by ``build.py`` with nvcc for sm_100a).

The library is a plain C ABI (``include/fedb200.h``): kernels take raw device pointers and a stream, so
there is no libtorch dependency and the build takes seconds.  On a machine with a GPU a missing or
unloadable library is an error — there is no silent PyTorch fallback for the fused paths.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_LIB: Optional[C.CDLL] = None
LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "lib", "libfedb200.so")

vp, i32, u32, u64, f32, f64, sz = C.c_void_p, C.c_int, C.c_uint, C.c_ulonglong, C.c_float, C.c_double, C.c_size_t

_PROTOS = {
    "fedb200_last_error": (C.c_char_p, []),
    "fedb200_abi_version": (i32, []),
    "fedb200_sm_count": (i32, []),
    "fedb200_set_pdl": (i32, [i32]),
    "fedb200_set_deterministic": (i32, [vp, sz]),
    "fedb200_get_deterministic": (i32, []),
    "fedb200_adamw_flat": (i32, [vp, vp, vp, vp, vp, sz, i32, f32, f32, f32, f32, f32, f32, i32, vp]),
    "fedb200_mix_n": (i32, [vp, vp, i32, vp, vp, sz, vp]),
    "fedb200_colsum_bf16": (i32, [vp, vp, i32, i32, i32, vp]),
    "fedb200_cast_f32_bf16": (i32, [vp, vp, sz, vp]),
    "fedb200_delta_bf16": (i32, [vp, vp, vp, sz, vp]),
    "fedb200_gemm_bf16": (i32, [i32, vp, i32, vp, i32, i32, i32, i32, i32, vp, vp, i32, vp, vp, i32, i32, u64, f32,
                                i32, i32, vp]),
    "fedb200_set_gemm_mode": (i32, [i32]),
    "fedb200_layernorm_fwd": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, f32, vp]),
    "fedb200_layernorm_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, u64, f32, vp]),
    "fedb200_embed_fwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, f32, u64, f32, vp]),
    "fedb200_embed_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, u64, f32, vp]),
    "fedb200_head_fwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, f32, u64, f32, vp]),
    "fedb200_head_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, u64, f32, vp]),
    "fedb200_attn_set_trace": (i32, [vp]),
    "fedb200_attn_set_diag": (i32, [i32, i32]),
    "fedb200_attn_fwd": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, f32, u64, f32, vp, vp]),
    "fedb200_attn_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, u64, f32, vp, vp]),
    "fedb200_gram_bf16": (i32, [vp, i32, i32, sz, sz, sz, sz, vp, i32, vp]),
    "fedb200_gram_bf16_seg": (i32, [vp, i32, i32, sz, i32, i32, i32, vp, i32, vp]),
    "fedb200_pagerank": (i32, [vp, i32, i32, f64, f64, i32, f64, i32, vp, vp, vp, vp]),
    "fedb200_sha256_leaves": (i32, [vp, vp, sz, u32, vp, vp]),
    "fedb200_merkle_root": (i32, [vp, vp, u32, vp, vp]),
    "fedb200_block_validate": (i32, [vp, vp, u32, vp, vp, vp]),
    "fedb200_sha256_root_static": (i32, [vp, sz, u32, vp, vp]),
    "fedb200_bert_forward": (i32, [vp, i32, i32, i32, i32, u64, i32, vp]),
    "fedb200_bert_backward": (i32, [vp, i32, i32, i32, u64, i32, vp]),
    "fedb200_peer_alloc": (i32, [sz, vp, vp]),
    "fedb200_peer_open": (i32, [vp, vp]),
    "fedb200_peer_close": (i32, [vp]),
    "fedb200_peer_free": (i32, [vp]),
    "fedb200_fedavg_reduce": (i32, [vp, vp, vp, vp, i32, i32, sz, f32, f32, u32, i32, vp]),
    "fedb200_gram_fused": (i32, [vp, i32, i32, sz, sz, sz, sz, vp, vp, vp, i32, u32, i32, i32, vp]),
    "fedb200_gram_fused_seg": (i32, [vp, i32, i32, sz, i32, i32, i32, vp, vp, vp, i32, u32, i32, i32, vp]),
    "fedb200_gossip_mix": (i32, [vp, vp, i32, vp, i32, vp, vp, i32, i32, vp, i32, i32, sz, u32, i32, vp]),
    "fedb200_gossip_mix_async": (i32, [vp, vp, vp, vp, vp, vp, i32, vp, vp, vp, vp, i32, sz, sz, u32, u32, vp, i32, i32, vp]),
    "fedb200_copy_flagged": (i32, [vp, vp, sz, sz, vp, u32, vp]),
    "fedb200_set_flags": (i32, [vp, u32, i32, vp]),
    "fedb200_fedavg_reduce_ex": (i32, [vp, vp, vp, vp, i32, i32, sz, f32, f32, u32, i32, i32, i32, vp]),
    "fedb200_fedavg_reduce_nvls": (i32, [vp, vp, vp, vp, vp, i32, i32, sz, f32, f32, u32, i32, i32, vp]),
    "fedb200_peer_exchange": (i32, [vp, vp, i32, i32, i32, u32, i32, vp]),
    "fedb200_peer_barrier": (i32, [vp, i32, i32, u32, i32, vp]),
    "fedb200_flag_set": (i32, [vp, u32, vp]),
    "fedb200_flag_wait": (i32, [vp, u32, i32, vp, vp]),
}


class BertLayer(C.Structure):
    """Mirror of ``fedb200_bert_layer`` (include/fedb200.h)."""
    _fields_ = [(n, vp) for n in (
        "wqkv", "wo", "w1", "w2", "bqkv", "bo", "b1", "b2", "ln1w", "ln1b", "ln2w", "ln2b",
        "g_wqkv", "g_bqkv", "g_wo", "g_bo", "g_w1", "g_b1", "g_w2", "g_b2", "g_ln1w", "g_ln1b", "g_ln2w", "g_ln2b",
        "x_in", "x_out", "qkv", "ctx", "z1", "a1", "u", "g", "z2", "mean1", "rstd1", "mean2", "rstd2", "lse", "drop_bits")]


MAX_LAYERS = 48


class BertPlan(C.Structure):
    """Mirror of ``fedb200_bert_plan`` (include/fedb200.h)."""
    _fields_ = ([(n, i32) for n in ("H", "I", "L", "heads", "C", "type_vocab", "act", "E")]
                + [(n, f32) for n in ("ln_eps", "p_hidden", "p_attn", "p_embed", "p_head")]
                + [(n, vp) for n in (
                    "word", "pos", "type", "emb_lnw", "emb_lnb", "Wp", "bp", "Wc", "bc",
                    "g_word", "g_pos", "g_type", "g_emb_lnw", "g_emb_lnb", "g_Wp", "g_bp", "g_Wc", "g_bc",
                    "emb_mean", "emb_rstd", "map_w", "map_b", "g_map_w", "g_map_b", "emb_out", "d_emb",
                    "ids", "pos_ids", "tt", "cu_seqlens", "labels",
                    "pooled", "pooled_drop", "dpre", "logits", "dlogits", "loss", "pred",
                    "dx", "dzb", "dfb", "dctx", "du", "dqkv", "dq_accum", "delta")]
                + [("layer", BertLayer * MAX_LAYERS)])


class KernelError(RuntimeError):
    pass


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        path = os.path.abspath(LIB_PATH)
        if not os.path.exists(path):
            raise KernelError(f"{path} is missing — run `python build.py` (nvcc, sm_100a) before using the fused ops")
        l = C.CDLL(path)
        for name, (res, args) in _PROTOS.items():
            if hasattr(l, name):
                fn = getattr(l, name)
                fn.restype, fn.argtypes = res, args
        _LIB = l
    return _LIB


def call(name: str, *args) -> None:
    l = lib()
    fn = getattr(l, name, None)
    if fn is None:
        raise KernelError(f"{name} is not exported by libfedb200.so (stale build?)")
    rc = fn(*args)
    if rc != 0:
        raise KernelError(f"{name} failed (cuda error {rc}): {l.fedb200_last_error().decode()}")


# number of kernel launches issued through this module (the bench reports it as gpu_launches)
LAUNCHES = 0


def count(n: int = 1) -> None:
    global LAUNCHES
    LAUNCHES += n
