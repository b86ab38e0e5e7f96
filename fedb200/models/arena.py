"""Flat parameter arena: every HF-keyed tensor is a view into ONE contiguous buffer.

The reference moves 201 separate numpy arrays per client per round
(``get_parameters`` src/Servercase/server_IID_IMDB.py:161-162) and averages them one by one
(src/Serverlesscase/serverless_IID_IMDB.py:269).  Here a whole model is a single
16-byte-aligned span so aggregation, hashing, scoring and AdamW are one launch over
``D`` elements (108.3 M for BERT-base/BioBERT, 335 M for BERT-large).
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Iterable, Tuple

import torch

ALIGN_ELEMS = 64   # 256 B in fp32 / 128 B in bf16: keeps every view TMA- and float4-aligned


def param_layout(shapes: "OrderedDict[str, Tuple[int, ...]]") -> Tuple["OrderedDict[str, Tuple[int, int]]", int]:
    """name -> (offset, numel) with each tensor start rounded up to ``ALIGN_ELEMS``."""
    off, out = 0, OrderedDict()
    for name, shp in shapes.items():
        n = 1
        for s in shp:
            n *= int(s)
        out[name] = (off, n)
        off += (n + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS
    return out, off


class FlatArena:
    def __init__(self, shapes: "OrderedDict[str, Tuple[int, ...]]", dtype=torch.float32, device="cpu",
                 buffer: torch.Tensor | None = None, key_order: Iterable[str] | None = None):
        """``shapes`` order is the memory layout; ``key_order`` (default: same) is the order in which
        ``state_dict()`` lists the tensors — the HF key order, whatever the layout."""
        self.shapes = OrderedDict((k, tuple(v)) for k, v in shapes.items())
        self.key_order = list(key_order) if key_order is not None else list(self.shapes)
        assert sorted(self.key_order) == sorted(self.shapes)
        self.layout, self.numel = param_layout(self.shapes)
        if buffer is None:
            buffer = torch.zeros(self.numel, dtype=dtype, device=device)
        assert buffer.numel() >= self.numel and buffer.is_contiguous()
        self.flat = buffer[: self.numel] if buffer.numel() != self.numel else buffer
        self.views: "OrderedDict[str, torch.Tensor]" = OrderedDict(
            (k, self.flat[o:o + n].view(self.shapes[k])) for k, (o, n) in self.layout.items())

    def like(self, dtype=None, device=None) -> "FlatArena":
        return FlatArena(self.shapes, dtype or self.flat.dtype, device or self.flat.device, key_order=self.key_order)

    def payload_numel(self) -> int:
        return sum(n for _, n in self.layout.values())

    def load(self, state: Dict[str, torch.Tensor]) -> None:
        for k, v in self.views.items():
            v.copy_(state[k].to(v.dtype))

    def state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((k, self.views[k]) for k in self.key_order)

    @staticmethod
    def adopt(module: torch.nn.Module, device=None) -> "FlatArena":
        """Move a module's parameters into a fresh arena; parameters become views of it."""
        sd = OrderedDict((k, tuple(v.shape)) for k, v in module.state_dict().items())
        dev = device or next(module.parameters()).device
        arena = FlatArena(sd, torch.float32, dev)
        arena.load(module.state_dict())
        named = dict(module.named_parameters())
        named.update(dict(module.named_buffers()))
        for k, view in arena.views.items():
            named[k].data = view
        return arena
