"""Model factory — the ``AutoModelForSequenceClassification.from_pretrained`` stand-in
(src/Servercase/server_IID_IMDB.py:142-144; ``ignore_mismatched_sizes`` variant at
src/Servercase/server_noniid_medical_transcriptions.py:146-148)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch

from ..config import ModelConfig


@dataclass
class SequenceClassifierOutput:
    loss: Optional[torch.Tensor]
    logits: torch.Tensor


def build_model(config: ModelConfig, device="cpu", seed: Optional[int] = None,
                checkpoint: Optional[str] = None, ignore_mismatched_sizes: bool = False) -> torch.nn.Module:
    from .albert import AlbertForSequenceClassification
    from .bert import BertForSequenceClassification
    if seed is not None:
        torch.manual_seed(seed)
    cls = {"bert": BertForSequenceClassification, "albert": AlbertForSequenceClassification}[config.arch]
    model = cls(config)
    if checkpoint:
        from .hf_io import load_pretrained
        model.load_report = load_pretrained(model, checkpoint, ignore_mismatched_sizes=ignore_mismatched_sizes)["load_report"]
    return model.to(device)
