"""HF ``save_pretrained`` / ``from_pretrained`` compatible checkpoint I/O (C14).

The reference writes ``global_model.save_pretrained(dir)`` every round, overwriting
(src/Serverlesscase/serverless_IID_IMDB.py:280), then reports the on-disk size in GiB
(src/Serverlesscase/serverless_NonIID_IMDB.py:305-318).  It has no resume path.  Here the
layout is the same — ``config.json`` + ``model.safetensors`` keyed by HF state_dict names —
so checkpoints load in ``transformers``; ``fed_state.json`` adds round / ledger metadata for
``--resume``.
"""
from __future__ import annotations

import json
import os
from collections import OrderedDict
from typing import Dict, Optional

import torch

from ..config import ModelConfig


def hf_config_dict(c: ModelConfig) -> dict:
    base = dict(vocab_size=c.vocab_size, hidden_size=c.hidden_size, num_hidden_layers=c.num_layers,
                num_attention_heads=c.num_heads, intermediate_size=c.intermediate_size,
                max_position_embeddings=c.max_position, type_vocab_size=c.type_vocab_size,
                layer_norm_eps=c.layer_norm_eps, initializer_range=c.initializer_range,
                id2label={str(i): f"LABEL_{i}" for i in range(c.num_labels)},
                label2id={f"LABEL_{i}": i for i in range(c.num_labels)}, torch_dtype="float32")
    if c.arch == "bert":
        base.update(model_type="bert", architectures=["BertForSequenceClassification"], hidden_act="gelu",
                    hidden_dropout_prob=c.hidden_dropout, attention_probs_dropout_prob=c.attention_dropout,
                    pad_token_id=0, position_embedding_type="absolute")
    else:
        base.update(model_type="albert", architectures=["AlbertForSequenceClassification"],
                    embedding_size=c.embedding_size, hidden_act="gelu_new", num_hidden_groups=1, inner_group_num=1,
                    hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, classifier_dropout_prob=0.1)
    return base


def config_from_hf(d: dict, name: str = "loaded") -> ModelConfig:
    arch = d.get("model_type", "bert")
    return ModelConfig(arch=arch, vocab_size=d["vocab_size"], hidden_size=d["hidden_size"],
                       num_layers=d["num_hidden_layers"], num_heads=d["num_attention_heads"],
                       intermediate_size=d["intermediate_size"], max_position=d["max_position_embeddings"],
                       type_vocab_size=d.get("type_vocab_size", 2), embedding_size=d.get("embedding_size", 128),
                       num_labels=len(d.get("id2label", {0: 0, 1: 1})), layer_norm_eps=d.get("layer_norm_eps", 1e-12),
                       hidden_dropout=d.get("hidden_dropout_prob", 0.1),
                       attention_dropout=d.get("attention_probs_dropout_prob", 0.1), name=name)


def _save_tensors(state: Dict[str, torch.Tensor], path: str) -> str:
    cpu = OrderedDict((k, v.detach().to("cpu", torch.float32).contiguous().clone()) for k, v in state.items())
    try:
        from safetensors.torch import save_file
        fn = os.path.join(path, "model.safetensors")
        save_file(cpu, fn, metadata={"format": "pt"})
    except ImportError:                      # pragma: no cover - safetensors ships with transformers
        fn = os.path.join(path, "pytorch_model.bin")
        torch.save(cpu, fn)
    return fn


def save_pretrained(state: Dict[str, torch.Tensor], config: ModelConfig, path: str,
                    fed_state: Optional[dict] = None) -> float:
    """Write an HF-layout checkpoint directory; returns its size in GiB."""
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "config.json"), "w") as fh:
        json.dump(hf_config_dict(config), fh, indent=2)
    _save_tensors(state, path)
    if fed_state is not None:
        with open(os.path.join(path, "fed_state.json"), "w") as fh:
            json.dump(fed_state, fh, indent=1, default=str)
    return dir_size_gib(path)


def load_state(path: str) -> "OrderedDict[str, torch.Tensor]":
    st = os.path.join(path, "model.safetensors")
    if os.path.exists(st):
        from safetensors.torch import load_file
        return OrderedDict(load_file(st))
    return OrderedDict(torch.load(os.path.join(path, "pytorch_model.bin"), map_location="cpu"))


def load_pretrained(model: torch.nn.Module, path: str, ignore_mismatched_sizes: bool = False) -> dict:
    """Load an HF-layout checkpoint into ``model``; returns ``fed_state`` (``{}`` if absent)."""
    state = load_state(path)
    own = model.state_dict()
    if ignore_mismatched_sizes:
        state = OrderedDict((k, v) for k, v in state.items() if k in own and own[k].shape == v.shape)
        own.update(state)
        state = own
    model.load_state_dict(state, strict=True)
    fs = os.path.join(path, "fed_state.json")
    if os.path.exists(fs):
        with open(fs) as fh:
            return json.load(fh)
    return {}


def load_config(path: str) -> ModelConfig:
    with open(os.path.join(path, "config.json")) as fh:
        return config_from_hf(json.load(fh))


def dir_size_gib(path: str) -> float:
    """``os.walk`` sum in GiB (src/Serverlesscase/serverless_NonIID_IMDB.py:307-318)."""
    total = 0
    for root, _, files in os.walk(path):
        for f in files:
            total += os.path.getsize(os.path.join(root, f))
    return total / (1024 ** 3)
