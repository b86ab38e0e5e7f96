"""Deterministic fault injector (new; SURVEY §5.3 — the reference has no fault injection, the
"2 injected anomalous nodes" config comes from BASELINE.json).  Corrupts a peer's *update*
``w_k - w_global``: scaled, sign-flipped or replaced by noise."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence

import numpy as np
import torch


def inject_anomaly(flat: torch.Tensor, reference: torch.Tensor, kind: str = "scale", strength: float = 10.0,
                   seed: int = 0) -> None:
    """In place: ``flat <- reference + corrupt(flat - reference)``."""
    delta = flat - reference
    if kind == "scale":
        delta.mul_(strength)
    elif kind == "signflip":
        delta.mul_(-strength)
    elif kind == "noise":
        g = torch.Generator(device="cpu").manual_seed(seed)
        rms = float(delta.pow(2).mean().sqrt()) or 1e-3
        noise = torch.randn(delta.numel(), generator=g, dtype=torch.float32).to(delta.device)
        delta.copy_(noise.view_as(delta) * rms * strength)
    else:
        raise ValueError(kind)
    flat.copy_(reference + delta)


@dataclass
class AnomalyInjector:
    num_peers: int
    count: int = 0
    kind: str = "scale"
    strength: float = 10.0
    seed: int = 0

    def targets(self) -> List[int]:
        if self.count <= 0:
            return []
        rng = np.random.default_rng(self.seed + 99991)
        return sorted(int(x) for x in rng.choice(self.num_peers, size=min(self.count, self.num_peers), replace=False))

    def maybe_corrupt(self, peer: int, flat: torch.Tensor, reference: torch.Tensor, round_idx: int) -> bool:
        if peer in self.targets():
            inject_anomaly(flat, reference, self.kind, self.strength, self.seed + 1000 * round_idx + peer)
            return True
        return False
