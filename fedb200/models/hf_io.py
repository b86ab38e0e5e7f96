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
    try:                                     # written under a temporary name and renamed: a reader never sees half a file
        from safetensors.torch import save_file
        fn = os.path.join(path, "model.safetensors")
        save_file(cpu, fn + ".tmp", metadata={"format": "pt"})
    except ImportError:                      # pragma: no cover - safetensors ships with transformers
        fn = os.path.join(path, "pytorch_model.bin")
        torch.save(cpu, fn + ".tmp")
    os.replace(fn + ".tmp", fn)
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


def _canonical_keys(state: "OrderedDict[str, torch.Tensor]", own_keys, arch: str) -> "OrderedDict[str, torch.Tensor]":
    """Key fix-ups ``transformers`` applies on load: old TF-style LayerNorm names (``gamma`` / ``beta``, e.g. the original
    BioBERT conversion) and checkpoints of the bare encoder (``embeddings.…`` without the ``bert.`` / ``albert.`` prefix)."""
    out = OrderedDict()
    prefix = arch + "."
    for k, v in state.items():
        if k.endswith("LayerNorm.gamma"):
            k = k[: -len("gamma")] + "weight"
        elif k.endswith("LayerNorm.beta"):
            k = k[: -len("beta")] + "bias"
        if k not in own_keys and (prefix + k) in own_keys:
            k = prefix + k
        out[k] = v
    return out


def load_pretrained(model: torch.nn.Module, path: str, ignore_mismatched_sizes: bool = False,
                    strict: bool = False) -> dict:
    """Load an HF-layout checkpoint directory into ``model`` with ``from_pretrained`` semantics
    (src/Servercase/server_IID_IMDB.py:142-144): checkpoint keys the model does not have (e.g. the ``cls.predictions.*``
    pre-training head of BioBERT) are ignored, model keys the checkpoint lacks (the fresh ``classifier``) keep their
    initialisation, a shape mismatch is an error unless ``ignore_mismatched_sizes`` (then that tensor also keeps its
    initialisation — src/Servercase/server_noniid_medical_transcriptions.py:146-148).  ``strict=True`` demands an exact key
    match (our own checkpoints, ``--resume``).  Returns ``fed_state`` (``{}`` if absent) with a ``"load_report"`` entry."""
    own = model.state_dict()
    arch = "albert" if any(k.startswith("albert.") for k in own) else "bert"
    state = _canonical_keys(load_state(path), own, arch)
    unexpected = [k for k in state if k not in own]
    missing = [k for k in own if k not in state]
    mismatched = [k for k in state if k in own and tuple(own[k].shape) != tuple(state[k].shape)]
    if strict and (unexpected or missing):
        raise KeyError(f"checkpoint key mismatch (strict): missing {missing[:3]}, unexpected {unexpected[:3]}")
    if mismatched and not ignore_mismatched_sizes:
        k = mismatched[0]
        raise RuntimeError(f"size mismatch for {k}: checkpoint {tuple(state[k].shape)} vs model {tuple(own[k].shape)} "
                           "(pass ignore_mismatched_sizes=True / --ignore-mismatched-sizes to re-initialise it)")
    for k, v in state.items():
        if k in own and k not in mismatched:
            own[k] = v
    model.load_state_dict(own, strict=True)
    out = {}
    fs = os.path.join(path, "fed_state.json")
    if os.path.exists(fs):
        with open(fs) as fh:
            out = json.load(fh)
    out["load_report"] = {"missing": missing, "unexpected": unexpected, "mismatched": mismatched}
    return out


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
