"""Fused BERT trainer on the hand-written sm_100a kernels (kernels K8-K13).

Same model as :class:`fedb200.models.bert.BertForSequenceClassification` (the reference's
``AutoModelForSequenceClassification`` BERT/BioBERT, src/Servercase/server_IID_IMDB.py:142-144) and the
same training step as the reference's ``train`` loop (:108-118: forward, ``loss.backward()``,
``AdamW.step()``, ``zero_grad()``), but re-designed for the B200:

* parameters, gradients and Adam state are flat fp32 arenas (HF-keyed views) with a bf16 shadow of the
  weights for the tensor cores; the memory layout puts query/key/value weights back to back so the three
  projections are ONE ``[3H, H]`` GEMM (forward, dgrad and wgrad);
* tokens are packed (no pad tokens): the padding mask of the reference becomes per-sequence lengths inside
  the attention kernel, so GEMM/LayerNorm work scales with real tokens only;
* forward and backward are explicit sequences of kernel launches (no autograd graph): every GEMM runs on
  tcgen05 with its bias / dropout / residual / GELU / GELU' fused into the epilogue, weight gradients are
  split-K tensor-core GEMMs reduced with fp32 atomics straight into the gradient arena, LayerNorm backward
  also emits the neighbouring dense layer's bias gradient, dropout masks are regenerated from a counter RNG;
* AdamW (HF semantics) is one pass over the arena that also refreshes the bf16 shadow and zeroes the grads.
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from ..config import FedConfig, ModelConfig
from ..data.dataset import collate_packed
from ..trainer.adamw import FlatAdamW
from .arena import FlatArena
from .factory import build_model


def _layer_keys(l: int) -> Dict[str, str]:
    p = f"bert.encoder.layer.{l}."
    return {
        "q.w": p + "attention.self.query.weight", "k.w": p + "attention.self.key.weight",
        "v.w": p + "attention.self.value.weight", "q.b": p + "attention.self.query.bias",
        "k.b": p + "attention.self.key.bias", "v.b": p + "attention.self.value.bias",
        "o.w": p + "attention.output.dense.weight", "o.b": p + "attention.output.dense.bias",
        "ln1.w": p + "attention.output.LayerNorm.weight", "ln1.b": p + "attention.output.LayerNorm.bias",
        "f1.w": p + "intermediate.dense.weight", "f1.b": p + "intermediate.dense.bias",
        "f2.w": p + "output.dense.weight", "f2.b": p + "output.dense.bias",
        "ln2.w": p + "output.LayerNorm.weight", "ln2.b": p + "output.LayerNorm.bias",
    }


def _albert_layer_keys() -> Dict[str, str]:
    p = "albert.encoder.albert_layer_groups.0.albert_layers.0."
    return {
        "q.w": p + "attention.query.weight", "k.w": p + "attention.key.weight", "v.w": p + "attention.value.weight",
        "q.b": p + "attention.query.bias", "k.b": p + "attention.key.bias", "v.b": p + "attention.value.bias",
        "o.w": p + "attention.dense.weight", "o.b": p + "attention.dense.bias",
        "ln1.w": p + "attention.LayerNorm.weight", "ln1.b": p + "attention.LayerNorm.bias",
        "f1.w": p + "ffn.weight", "f1.b": p + "ffn.bias", "f2.w": p + "ffn_output.weight", "f2.b": p + "ffn_output.bias",
        "ln2.w": p + "full_layer_layer_norm.weight", "ln2.b": p + "full_layer_layer_norm.bias",
    }


_LAYER_ORDER = ("q.w", "k.w", "v.w", "q.b", "k.b", "v.b", "o.w", "o.b", "ln1.w", "ln1.b", "f1.w", "f1.b", "f2.w", "f2.b",
                "ln2.w", "ln2.b")


def _roles(cfg: ModelConfig):
    """HF key names of every tensor the engine touches, per architecture.  ALBERT shares ONE layer's weights
    across all ``num_layers`` applications (every entry of ``layers`` is the same dict) and adds the
    128 -> H embedding projection."""
    if cfg.arch == "bert":
        e = "bert.embeddings."
        return dict(word=e + "word_embeddings.weight", pos=e + "position_embeddings.weight",
                    type=e + "token_type_embeddings.weight", emb_lnw=e + "LayerNorm.weight", emb_lnb=e + "LayerNorm.bias",
                    map_w=None, map_b=None, layers=[_layer_keys(l) for l in range(cfg.num_layers)],
                    Wp="bert.pooler.dense.weight", bp="bert.pooler.dense.bias", Wc="classifier.weight", bc="classifier.bias")
    e = "albert.embeddings."
    lk = _albert_layer_keys()
    return dict(word=e + "word_embeddings.weight", pos=e + "position_embeddings.weight",
                type=e + "token_type_embeddings.weight", emb_lnw=e + "LayerNorm.weight", emb_lnb=e + "LayerNorm.bias",
                map_w="albert.encoder.embedding_hidden_mapping_in.weight", map_b="albert.encoder.embedding_hidden_mapping_in.bias",
                layers=[lk] * cfg.num_layers, Wp="albert.pooler.weight", bp="albert.pooler.bias", Wc="classifier.weight",
                bc="classifier.bias")


class FusedRunner:
    """Drop-in for :class:`fedb200.trainer.local.TorchRunner` on a CUDA device (BERT and ALBERT families)."""

    backend = "fused"

    def __init__(self, model_cfg: ModelConfig, fed: Optional[FedConfig] = None, device="cuda", seed: Optional[int] = 0,
                 max_tokens: Optional[int] = None, max_batch: Optional[int] = None):
        if model_cfg.arch not in ("bert", "albert"):
            raise ValueError(f"no fused implementation for architecture {model_cfg.arch!r}")
        if model_cfg.head_dim != 64:
            raise ValueError("fused attention is specialised for head_dim 64 (BERT-base/large, BioBERT)")
        from ..ops import cuda_ops           # fails loudly if libfedb200.so is missing
        self.ops = cuda_ops
        fed = fed or FedConfig()
        self.cfg, self.fed, self.device = model_cfg, fed, torch.device(device)
        c = model_cfg
        self.H, self.I, self.L, self.heads, self.C = c.hidden_size, c.intermediate_size, c.num_layers, c.num_heads, c.num_labels
        self.train_batch = fed.data.batch_size
        # forward-only passes may pack several reference batches into one launch sequence (same samples, same loss
        # normalisation, bigger GEMMs): the sequence-count capacity is 4 reference batches, the token capacity is
        # what one training batch can need
        self.max_batch = max_batch or 4 * fed.data.batch_size
        self.max_tokens = max_tokens or fed.data.batch_size * min(fed.data.max_seq_len, c.max_position)

        # ---- arenas: HF keys, compute-friendly memory order (q|k|v weights contiguous) -----------------
        ref = build_model(c, "cpu", seed, checkpoint=fed.init_checkpoint, ignore_mismatched_sizes=fed.ignore_mismatched_sizes)
        self.load_report = getattr(ref, "load_report", None)
        sd = ref.state_dict()
        hf_order = list(sd.keys())
        self.roles = R = _roles(c)
        self.E = c.embedding_size if c.arch == "albert" else c.hidden_size
        order = [R["word"], R["pos"], R["type"], R["emb_lnw"], R["emb_lnb"]]
        if R["map_w"]:
            order += [R["map_w"], R["map_b"]]
        self.lk = R["layers"]
        seen = set()
        for k in self.lk:
            if id(k) in seen:
                continue
            seen.add(id(k))
            order += [k[n] for n in _LAYER_ORDER]
        order += [R["Wp"], R["bp"], R["Wc"], R["bc"]]
        assert sorted(order) == sorted(hf_order), "unexpected parameter set"
        shapes = OrderedDict((k, tuple(sd[k].shape)) for k in order)
        self.arena = FlatArena(shapes, torch.float32, self.device, key_order=hf_order)
        self.arena.load(sd)
        del ref, sd
        self.grads = self.arena.like()
        self.shadow = self.arena.like(dtype=torch.bfloat16)
        self.opt = FlatAdamW(self.arena.flat, self.grads.flat, fed.lr, fed.adam_betas, fed.adam_eps, fed.weight_decay,
                             bf16_shadow=self.shadow.flat, use_kernel=True)
        self.sync_shadow()
        H = self.H
        # fused [3H, H] / [3H] views over the contiguous q|k|v spans
        def fused(arena: FlatArena, l: int):
            k = self.lk[l]
            ow, _ = arena.layout[k["q.w"]]
            ob, _ = arena.layout[k["q.b"]]
            assert arena.layout[k["k.w"]][0] == ow + H * H and arena.layout[k["v.w"]][0] == ow + 2 * H * H
            assert arena.layout[k["k.b"]][0] == ob + H and arena.layout[k["v.b"]][0] == ob + 2 * H
            return arena.flat[ow:ow + 3 * H * H].view(3 * H, H), arena.flat[ob:ob + 3 * H]
        self.wqkv_bf = [fused(self.shadow, l)[0] for l in range(self.L)]
        self.bqkv = [fused(self.arena, l)[1] for l in range(self.L)]
        self.gwqkv = [fused(self.grads, l)[0] for l in range(self.L)]
        self.gbqkv = [fused(self.grads, l)[1] for l in range(self.L)]

        # ---- activation workspace (sized once for max_tokens) ---------------------------------------------
        T, Bm, I, L, dev = self.max_tokens, self.max_batch, self.I, self.L, self.device
        bf = lambda *s: torch.empty(*s, dtype=torch.bfloat16, device=dev)
        f32 = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
        self.xs = [bf(T, H) for _ in range(L + 1)]
        self.act = [dict(qkv=bf(T, 3 * H), ctx=bf(T, H), z1=bf(T, H), a1=bf(T, H), u=bf(T, I), g=bf(T, I), z2=bf(T, H),
                         mean1=f32(T), rstd1=f32(T), mean2=f32(T), rstd2=f32(T), lse=f32(self.heads * T),
                         drop_bits=(torch.empty(self.heads * T * ((c.max_position + 31) // 32), dtype=torch.int32, device=dev)
                                    if (c.attention_dropout > 0 and c.arch == "bert") else None))
                    for _ in range(L)]
        self.emb_mean, self.emb_rstd = f32(T), f32(T)
        self.emb_out = bf(T, self.E) if self.E != H else None
        self.d_emb = bf(T, self.E) if self.E != H else None
        self.dx, self.dzb, self.dfb, self.dctx = bf(T, H), bf(T, H), bf(T, H), bf(T, H)
        self.du, self.dqkv = bf(T, I), bf(T, 3 * H)
        self.dq_accum = torch.empty((c.max_position + 127) // 128 * T, H, dtype=torch.float32, device=dev)   # dQ partial slots
        self.delta = f32(self.heads * T)
        self.pooled, self.pooled_drop, self.dpre = f32(Bm, H), f32(Bm, H), f32(Bm, H)
        self.logits, self.dlogits = f32(Bm, self.C), f32(Bm, self.C)
        self.loss_vec = f32(Bm)
        self.pred = torch.empty(Bm, dtype=torch.int32, device=dev)
        self.d_ids = torch.empty(T, dtype=torch.int32, device=dev)
        self.d_pos = torch.empty(T, dtype=torch.int32, device=dev)
        self.d_tt = torch.empty(T, dtype=torch.int32, device=dev)
        self.d_cu = torch.empty(Bm + 1, dtype=torch.int32, device=dev)
        self.d_labels = torch.empty(Bm, dtype=torch.int32, device=dev)
        self._dev = {"ids": self.d_ids, "pos": self.d_pos, "tt": self.d_tt, "cu_seqlens": self.d_cu, "labels": self.d_labels}
        pin = lambda n: torch.empty(n, dtype=torch.int32).pin_memory()
        self._stage = [{"ids": pin(T), "pos": pin(T), "tt": pin(T), "cu_seqlens": pin(Bm + 1), "labels": pin(Bm),
                        "event": None} for _ in range(2)]
        self._slot = 0
        self.step_count = 0
        self.samples_seen = 0
        self.tokens_seen = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.base_seed = (int(seed or 0) * 0x9E3779B1 + 0x1234567) & 0xFFFFFFFF
        self.use_engine = True                  # native step executor (csrc/bert_engine.cu); False = Python twin
        self._plan = self._build_plan()
        self._sms = torch.cuda.get_device_properties(self.device).multi_processor_count
        self.det_ws: Optional[torch.Tensor] = None
        if fed.deterministic:
            self.set_deterministic(True)

    def set_deterministic(self, on: bool, workspace_mb: int = 256) -> None:
        """Bit-reproducible training (``--deterministic``, csrc/determinism.cu): split-K weight gradients, bias / LayerNorm
        column sums and the embedding-table scatter store per-split / per-CTA partials in a workspace and add them in a
        fixed order instead of with fp32 atomics.  Two runs from the same seed, data and SM count then end with the same
        bytes, so a ledger's Merkle roots can be re-derived by an auditor.  The switch is process-wide (one trainer per
        process / GPU is the deployment model); the workspace belongs to this runner."""
        from ..ops import _lib
        if on:
            need = max(int(workspace_mb) << 20, self.max_tokens * max(self.E, self.H) * 4 + (64 << 20))
            self.det_ws = torch.empty(need, dtype=torch.uint8, device=self.device)
            _lib.call("fedb200_set_deterministic", self.det_ws.data_ptr(), need)
        else:
            torch.cuda.synchronize(self.device)      # nothing in flight may still use the workspace
            _lib.call("fedb200_set_deterministic", None, 0)
            self.det_ws = None

    def _build_plan(self):
        """Fill the C plan with raw device pointers (valid for the life of the runner: nothing is reallocated)."""
        import ctypes as C
        from ..ops._lib import BertPlan, MAX_LAYERS
        c = self.cfg
        if self.L > MAX_LAYERS:
            raise ValueError(f"{self.L} layers exceed the native plan capacity ({MAX_LAYERS})")
        P = BertPlan()
        R = self.roles
        albert = c.arch == "albert"
        P.H, P.I, P.L, P.heads, P.C, P.type_vocab = self.H, self.I, self.L, self.heads, self.C, c.type_vocab_size
        P.act, P.E = (1 if albert else 0), self.E
        P.ln_eps = c.layer_norm_eps
        if albert:       # albert-base-v2: no encoder dropout, classifier dropout 0.1 (see models/albert.py)
            P.p_hidden, P.p_attn, P.p_embed, P.p_head = 0.0, 0.0, 0.0, (0.1 if c.hidden_dropout > 0 else 0.0)
        else:
            P.p_hidden, P.p_attn, P.p_embed, P.p_head = c.hidden_dropout, c.attention_dropout, c.hidden_dropout, c.hidden_dropout
        w, g = (lambda k: self.arena.views[k].data_ptr()), (lambda k: self.grads.views[k].data_ptr())
        for f in ("word", "pos", "type", "emb_lnw", "emb_lnb", "Wp", "bp", "Wc", "bc"):
            setattr(P, f, w(R[f]))
            setattr(P, "g_" + f, g(R[f]))
        if albert:
            P.map_w, P.map_b = self.shadow.views[R["map_w"]].data_ptr(), w(R["map_b"])
            P.g_map_w, P.g_map_b = g(R["map_w"]), g(R["map_b"])
            P.emb_out, P.d_emb = self.emb_out.data_ptr(), self.d_emb.data_ptr()
        for f, t in (("emb_mean", self.emb_mean), ("emb_rstd", self.emb_rstd), ("ids", self.d_ids), ("pos_ids", self.d_pos),
                     ("tt", self.d_tt), ("cu_seqlens", self.d_cu), ("labels", self.d_labels), ("pooled", self.pooled),
                     ("pooled_drop", self.pooled_drop), ("dpre", self.dpre), ("logits", self.logits),
                     ("dlogits", self.dlogits), ("loss", self.loss_vec), ("pred", self.pred), ("dx", self.dx),
                     ("dzb", self.dzb), ("dfb", self.dfb), ("dctx", self.dctx), ("du", self.du), ("dqkv", self.dqkv),
                     ("dq_accum", self.dq_accum), ("delta", self.delta)):
            setattr(P, f, t.data_ptr())
        for l in range(self.L):
            k, a, Ly = self.lk[l], self.act[l], P.layer[l]
            Ly.wqkv, Ly.bqkv = self.wqkv_bf[l].data_ptr(), self.bqkv[l].data_ptr()
            Ly.g_wqkv, Ly.g_bqkv = self.gwqkv[l].data_ptr(), self.gbqkv[l].data_ptr()
            for f, key in (("wo", "o.w"), ("w1", "f1.w"), ("w2", "f2.w")):
                setattr(Ly, f, self.shadow.views[k[key]].data_ptr())
                setattr(Ly, "g_" + f, g(k[key]))
            for f, key in (("bo", "o.b"), ("b1", "f1.b"), ("b2", "f2.b"), ("ln1w", "ln1.w"), ("ln1b", "ln1.b"),
                           ("ln2w", "ln2.w"), ("ln2b", "ln2.b")):
                setattr(Ly, f, w(k[key]))
                setattr(Ly, "g_" + f, g(k[key]))
            Ly.x_in, Ly.x_out = self.xs[l].data_ptr(), self.xs[l + 1].data_ptr()
            for f in ("qkv", "ctx", "z1", "a1", "u", "g", "z2", "mean1", "rstd1", "mean2", "rstd2", "lse"):
                setattr(Ly, f, a[f].data_ptr())
            Ly.drop_bits = a["drop_bits"].data_ptr() if a["drop_bits"] is not None else None
        self._plan_ref = C.byref(P)
        return P

    # ---- parameters ----------------------------------------------------------------------------------------
    @property
    def flat(self) -> torch.Tensor:
        return self.arena.flat

    def state_dict(self):
        return self.arena.state_dict()

    def load_flat(self, flat: torch.Tensor) -> None:
        """Overwrite the fp32 master with ``flat``.  The copy also writes the bf16 shadow in the same pass (one read of
        the source instead of copy + cast); the ``sync_shadow()`` that callers issue right after is then a no-op."""
        if flat.data_ptr() != self.arena.flat.data_ptr():
            if flat.is_cuda and flat.dtype == torch.float32 and flat.is_contiguous() and flat.numel() == self.arena.flat.numel():
                self.ops.mix_n([flat], [1.0], self.arena.flat, self.shadow.flat)
                self._shadow_fresh = True
            else:
                self.arena.flat.copy_(flat)
                self._shadow_fresh = False

    def load_state_dict(self, state: Dict[str, torch.Tensor]) -> None:
        missing = set(self.arena.views) ^ set(state)
        if missing:
            raise KeyError(f"state_dict key mismatch (strict): {sorted(missing)[:4]} …")
        self.arena.load(state)
        self._shadow_fresh = False                  # the master changed behind the latch: the next sync_shadow() must cast

    def mark_dirty(self) -> None:
        """Call after editing ``flat`` in place (e.g. fault injection): the bf16 shadow no longer matches the master."""
        self._shadow_fresh = False

    def sync_shadow(self) -> None:
        """Refresh the bf16 tensor-core copy after the fp32 master was overwritten (load / aggregation)."""
        if getattr(self, "_shadow_fresh", False):      # load_flat() just wrote it; anything else since must call us again
            self._shadow_fresh = False
            return
        self.ops.cast_bf16(self.arena.flat, self.shadow.flat)

    # ---- helpers -----------------------------------------------------------------------------------------------
    def _w(self, key: str) -> torch.Tensor:
        return self.arena.views[key]

    def _wb(self, key: str) -> torch.Tensor:
        return self.shadow.views[key]

    def _g(self, key: str) -> torch.Tensor:
        return self.grads.views[key]

    def _seed(self, layer: int, which: int) -> int:
        return ((self.base_seed ^ (self.step_count * 0x85EBCA6B)) * 31 + layer * 7 + which) & 0x7FFFFFFFFFFF

    @staticmethod
    def _splits(out_rows: int, out_cols: int, tokens: int, sms: int = 148) -> int:
        """Split-K factor for a weight-gradient GEMM: the output has few tiles (18-72), so the token
        reduction is split until the persistent grid is filled with little wave-quantisation loss."""
        bn = 256 if out_cols >= 256 else (128 if out_cols >= 128 else 64)
        tiles = ((out_rows + 127) // 128) * ((out_cols + bn - 1) // bn)
        kmax = max(1, (tokens + 63) // 64)
        best, best_eff = 1, 0.0
        for s in range(1, min(32, kmax) + 1):
            total = tiles * s
            eff = total / (((total + sms - 1) // sms) * sms)
            if total >= int(0.9 * sms) and eff >= 0.9:
                return s
            if eff > best_eff:
                best, best_eff = s, eff
        return best

    def _upload(self, batch: Dict) -> Tuple[int, int, int]:
        """Host -> device copy of one packed batch through a double-buffered pinned staging area."""
        T, B = int(batch["num_tokens"]), int(batch["labels"].numel())
        if T > self.max_tokens or B > self.max_batch:
            raise ValueError(f"batch of {B} sequences / {T} tokens exceeds the workspace ({self.max_batch} / {self.max_tokens})")
        if batch["ids"].is_cuda:                 # pre-staged on the device (kernel-only timing): plain d2d copies
            for key, n in (("ids", T), ("pos", T), ("tt", T), ("cu_seqlens", B + 1), ("labels", B)):
                self._dev[key][:n].copy_(batch[key], non_blocking=True)
            return T, B, int(batch["max_len"])
        slot = self._slot
        self._slot ^= 1
        st = self._stage[slot]
        if st["event"] is not None:
            st["event"].synchronize()            # the copy issued from this slot two steps ago has landed
        for key, n in (("ids", T), ("pos", T), ("tt", T), ("cu_seqlens", B + 1), ("labels", B)):
            st[key][:n].copy_(batch[key])
            self._dev[key][:n].copy_(st[key][:n], non_blocking=True)
            self.h2d_bytes += n * 4
        if st["event"] is None:
            st["event"] = torch.cuda.Event()
        st["event"].record()
        return T, B, int(batch["max_len"])

    # ---- forward ---------------------------------------------------------------------------------------------------
    def _forward(self, T: int, B: int, max_len: int, train: bool) -> None:
        if not self.use_engine:
            return self._forward_py(T, B, max_len, train)
        from ..ops import _lib
        _lib.call("fedb200_bert_forward", self._plan_ref, T, B, max_len, int(train), self._step_seed(), self._sms,
                  torch.cuda.current_stream().cuda_stream)
        _lib.count(3 + 7 * self.L)

    def _backward(self, T: int, B: int, max_len: int) -> None:
        if not self.use_engine:
            return self._backward_py(T, B, max_len)
        from ..ops import _lib
        _lib.call("fedb200_bert_backward", self._plan_ref, T, B, max_len, self._step_seed(), self._sms,
                  torch.cuda.current_stream().cuda_stream)
        _lib.count(4 + 16 * self.L + ((4 + 8 * self.L) if self.det_ws is not None else 0))   # + the ordered-sum kernels

    def _step_seed(self) -> int:
        return (self.base_seed * 0x100000001B3 + self.step_count) & 0x7FFFFFFFFFFFFFFF

    def _forward_py(self, T: int, B: int, max_len: int, train: bool) -> None:
        """Python twin of csrc/bert_engine.cu::fedb200_bert_forward (kept for debugging / documentation; BERT only)."""
        if self.cfg.arch != "bert":
            raise NotImplementedError("the Python twin covers BERT; ALBERT runs on the native executor only")
        ops, H, I, c = self.ops, self.H, self.I, self.cfg
        ph = c.hidden_dropout if train else 0.0
        pa = c.attention_dropout if train else 0.0
        ids, pos, tt, cu = self.d_ids[:T], self.d_pos[:T], self.d_tt[:T], self.d_cu[:B + 1]
        ops.embed_fwd(ids, pos, tt, self._w("bert.embeddings.word_embeddings.weight"),
                      self._w("bert.embeddings.position_embeddings.weight"),
                      self._w("bert.embeddings.token_type_embeddings.weight"),
                      self._w("bert.embeddings.LayerNorm.weight"), self._w("bert.embeddings.LayerNorm.bias"),
                      self.xs[0][:T], self.emb_mean[:T], self.emb_rstd[:T], c.layer_norm_eps, self._seed(99, 0), ph)
        scale = 1.0 / math.sqrt(c.head_dim)
        for l in range(self.L):
            k, a = self.lk[l], self.act[l]
            x = self.xs[l][:T]
            ops.gemm(0, x, self.wqkv_bf[l], a["qkv"][:T], T, 3 * H, H, ops.EPI_BIAS, bias=self.bqkv[l])
            ops.attn_fwd(a["qkv"][:T], a["ctx"][:T], a["lse"], cu, B, self.heads, max_len, scale, self._seed(l, 1), pa,
                         drop_bits=a["drop_bits"] if train else None)
            ops.gemm(0, a["ctx"][:T], self._wb(k["o.w"]), a["z1"][:T], T, H, H, ops.EPI_BIAS_DROP_RES,
                     bias=self._w(k["o.b"]), aux=x, seed=self._seed(l, 2), dropout_p=ph)
            ops.layernorm_fwd(a["z1"][:T], self._w(k["ln1.w"]), self._w(k["ln1.b"]), a["a1"][:T], a["mean1"][:T],
                              a["rstd1"][:T], c.layer_norm_eps)
            ops.gemm(0, a["a1"][:T], self._wb(k["f1.w"]), a["u"][:T], T, I, H, ops.EPI_BIAS_GELU, bias=self._w(k["f1.b"]),
                     out2=a["g"][:T])
            ops.gemm(0, a["g"][:T], self._wb(k["f2.w"]), a["z2"][:T], T, H, I, ops.EPI_BIAS_DROP_RES,
                     bias=self._w(k["f2.b"]), aux=a["a1"][:T], seed=self._seed(l, 3), dropout_p=ph)
            ops.layernorm_fwd(a["z2"][:T], self._w(k["ln2.w"]), self._w(k["ln2.b"]), self.xs[l + 1][:T], a["mean2"][:T],
                              a["rstd2"][:T], c.layer_norm_eps)
        ops.head_fwd(self.xs[self.L][:T], cu, self._w("bert.pooler.dense.weight"), self._w("bert.pooler.dense.bias"),
                     self._w("classifier.weight"), self._w("classifier.bias"), self.d_labels[:B], self.pooled[:B],
                     self.pooled_drop[:B], self.logits[:B], self.loss_vec[:B], self.pred[:B], self.dlogits[:B], 1.0 / B,
                     self._seed(98, 0), ph)

    # ---- backward --------------------------------------------------------------------------------------------------
    def _backward_py(self, T: int, B: int, max_len: int) -> None:
        """Python twin of csrc/bert_engine.cu::fedb200_bert_backward."""
        ops, H, I, c = self.ops, self.H, self.I, self.cfg
        ph, pa = c.hidden_dropout, c.attention_dropout
        cu = self.d_cu[:B + 1]
        dx, dzb, dfb = self.dx[:T], self.dzb[:T], self.dfb[:T]
        df = dfb if ph > 0 else dzb
        dx.zero_()
        ops.head_bwd(self.xs[self.L][:T], cu, self._w("bert.pooler.dense.weight"), self._w("classifier.weight"),
                     self.pooled[:B], self.pooled_drop[:B], self.dlogits[:B], self.dpre[:B],
                     self._g("bert.pooler.dense.weight"), self._g("bert.pooler.dense.bias"), self._g("classifier.weight"),
                     self._g("classifier.bias"), dx, self._seed(98, 0), ph)
        scale = 1.0 / math.sqrt(c.head_dim)
        for l in reversed(range(self.L)):
            k, a = self.lk[l], self.act[l]
            x = self.xs[l][:T]
            # ---- FFN block ----
            ops.layernorm_bwd(dx, a["z2"][:T], a["mean2"][:T], a["rstd2"][:T], self._w(k["ln2.w"]), dzb, dfb,
                              self._g(k["ln2.w"]), self._g(k["ln2.b"]), self._g(k["f2.b"]), self._seed(l, 3), ph)
            ops.gemm(1, df, self._wb(k["f2.w"]), self.du[:T], T, I, H, ops.EPI_MUL_GELU_GRAD, aux=a["u"][:T])
            ops.gemm(2, df, a["g"][:T], self._g(k["f2.w"]), H, I, T, ops.EPI_ATOMIC_F32, splits=self._splits(H, I, T))
            ops.colsum_bf16(self.du[:T], self._g(k["f1.b"]))
            ops.gemm(1, self.du[:T], self._wb(k["f1.w"]), dx, T, H, I, ops.EPI_ADD, aux=dzb)
            ops.gemm(2, self.du[:T], a["a1"][:T], self._g(k["f1.w"]), I, H, T, ops.EPI_ATOMIC_F32, splits=self._splits(I, H, T))
            # ---- attention block ----
            ops.layernorm_bwd(dx, a["z1"][:T], a["mean1"][:T], a["rstd1"][:T], self._w(k["ln1.w"]), dzb, dfb,
                              self._g(k["ln1.w"]), self._g(k["ln1.b"]), self._g(k["o.b"]), self._seed(l, 2), ph)
            ops.gemm(1, df, self._wb(k["o.w"]), self.dctx[:T], T, H, H, ops.EPI_ADD)
            ops.gemm(2, df, a["ctx"][:T], self._g(k["o.w"]), H, H, T, ops.EPI_ATOMIC_F32, splits=self._splits(H, H, T))
            ops.attn_bwd(a["qkv"][:T], a["ctx"][:T], self.dctx[:T], a["lse"], self.delta, self.dqkv[:T],
                         self.dq_accum[:T], cu, B, self.heads, max_len, scale, self._seed(l, 1), pa, dbias=self.gbqkv[l],
                         drop_bits=a["drop_bits"])
            ops.gemm(1, self.dqkv[:T], self.wqkv_bf[l], dx, T, H, 3 * H, ops.EPI_ADD, aux=dzb)
            ops.gemm(2, self.dqkv[:T], x, self.gwqkv[l], 3 * H, H, T, ops.EPI_ATOMIC_F32, splits=self._splits(3 * H, H, T))
        ops.embed_bwd(dx, self.d_ids[:T], self.d_pos[:T], self.d_tt[:T], self._w("bert.embeddings.word_embeddings.weight"),
                      self._w("bert.embeddings.position_embeddings.weight"),
                      self._w("bert.embeddings.token_type_embeddings.weight"), self._w("bert.embeddings.LayerNorm.weight"),
                      self.emb_mean[:T], self.emb_rstd[:T], self._g("bert.embeddings.word_embeddings.weight"),
                      self._g("bert.embeddings.position_embeddings.weight"),
                      self._g("bert.embeddings.token_type_embeddings.weight"), self._g("bert.embeddings.LayerNorm.weight"),
                      self._g("bert.embeddings.LayerNorm.bias"), self._seed(99, 0), ph)

    # ---- public steps --------------------------------------------------------------------------------------------------
    def train_step(self, batch: Dict, optimizer_step: bool = True) -> torch.Tensor:
        """One forward + backward (+ AdamW) on a packed batch; returns the mean loss as a device scalar."""
        T, B, max_len = self._upload(batch)
        self._forward(T, B, max_len, train=True)
        self._backward(T, B, max_len)
        if optimizer_step:
            self.opt.step(zero_grad=True)
        self.step_count += 1
        self.samples_seen += B
        self.tokens_seen += T
        return self.loss_vec[:B].mean()

    def eval_step(self, batch: Dict) -> Tuple[torch.Tensor, torch.Tensor]:
        """Forward only; returns (mean loss, number of correct predictions) as device scalars."""
        T, B, max_len = self._upload(batch)
        self._forward(T, B, max_len, train=False)
        return self.loss_vec[:B].mean(), (self.pred[:B] == self.d_labels[:B]).sum()

    def forward_logits(self, batch: Dict) -> torch.Tensor:
        T, B, max_len = self._upload(batch)
        self._forward(T, B, max_len, train=False)
        return self.logits[:B].clone()

    # ---- runner interface (same as TorchRunner) ------------------------------------------------------------------------------
    def train_epochs(self, loader, epochs: int = 1, lazy: bool = False) -> Dict[str, float]:
        """``lazy``: return the last loss as a device scalar (the caller reads it back with the round's other
        results in one synchronisation, :class:`fedb200.utils.readback.HostReadback`)."""
        self.opt.reset()                          # fresh optimizer per call (server_IID_IMDB.py:109)
        loss, n = None, 0
        for _ in range(epochs):
            for idx in loader.batch_indices():
                loss = self.train_step(collate_packed(loader.dataset, idx, longest_first=True))
                n += 1
        if lazy:
            return {"loss": loss if loss is not None else torch.full((), float("nan"), device=self.device), "steps": n}
        out = float(loss.item()) if loss is not None else float("nan")     # one D2H read per local epoch
        self.d2h_bytes += 4
        return {"loss": out, "steps": n}

    def eval_plan(self, loader):
        """Pack the loader's (unshuffled) reference batches into as few forward passes as the workspace allows.
        Returns (list of packed micro-batches, list of reference batch sizes)."""
        ref = loader.batch_indices()
        lens = loader.dataset.lengths()
        micro, cur, cur_tok = [], [], 0
        for idx in ref:
            for i in idx:
                n = int(lens[i])
                if cur and (cur_tok + n > self.max_tokens or len(cur) >= self.max_batch):
                    micro.append(collate_packed(loader.dataset, np.asarray(cur), longest_first=True))
                    cur, cur_tok = [], 0
                cur.append(int(i))
                cur_tok += n
        if cur:
            micro.append(collate_packed(loader.dataset, np.asarray(cur), longest_first=True))
        return micro, [len(i) for i in ref]

    def evaluate_packed(self, micro, ref_sizes) -> Tuple[torch.Tensor, torch.Tensor]:
        """Forward passes over pre-packed micro-batches; returns device scalars (sum over REFERENCE batches of the
        batch-mean loss, number of correct predictions)."""
        losses, correct = [], torch.zeros((), dtype=torch.int64, device=self.device)
        for mb in micro:
            T, B, max_len = self._upload(mb)
            self._forward(T, B, max_len, train=False)
            lv = self.loss_vec[:B].clone()
            if mb.get("perm") is not None:                # packed longest-first: back to the loader's sample order
                back = torch.empty_like(lv)
                back[mb["perm"].to(self.device, non_blocking=True)] = lv
                lv = back
            losses.append(lv)
            correct += (self.pred[:B] == self.d_labels[:B]).sum()
        per_sample = torch.cat(losses)
        loss_sum = torch.zeros((), device=self.device)
        o = 0
        for n in ref_sizes:                       # mixed normalisation of the reference: sum of batch means
            loss_sum += per_sample[o:o + n].mean()
            o += n
        return loss_sum, correct

    def evaluate_lazy(self, loader):
        """Like :meth:`evaluate` without the host read: (loss_sum, correct) device scalars + (n_dataset, n_evaluated)."""
        micro, ref_sizes = self.eval_plan(loader)
        n_ds = max(len(loader.dataset), 1)
        if not micro:
            z = torch.zeros((), device=self.device)
            return z, z.to(torch.int64), n_ds, 0
        loss_sum, correct = self.evaluate_packed(micro, ref_sizes)
        return loss_sum, correct, n_ds, sum(ref_sizes)

    def evaluate(self, loader) -> Tuple[float, float]:
        micro, ref_sizes = self.eval_plan(loader)
        if not micro:
            return 0.0, 0.0
        loss_sum, correct = self.evaluate_packed(micro, ref_sizes)
        total = sum(ref_sizes)
        self.d2h_bytes += 12
        n_ds = max(len(loader.dataset), 1)
        # sum of batch-mean losses / len(dataset), as the reference computes it (server_IID_IMDB.py:133)
        return float(loss_sum.item()) / n_ds, (int(correct.item()) / total if total else 0.0)
