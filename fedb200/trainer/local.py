"""Local training / evaluation loops (C5, C6) behind one runner interface.

Reference loops: ``train(net, loader, epochs)`` src/Servercase/server_IID_IMDB.py:108-118 and
``test(net, loader)`` :121-135 (method copies ``train_model`` / ``evaluate_model`` at
src/Serverlesscase/serverless_IID_IMDB.py:156-187).  Quirks kept on purpose:

* a fresh optimizer per ``train`` call (state never persists across rounds, :109);
* ``test`` returns ``sum(batch-mean losses) / len(dataset)`` — mixed normalisation (:133).

``TorchRunner`` drives a plain ``nn.Module`` (CPU plumbing path and fp32 oracle);
:class:`fedb200.models.fused_bert.FusedRunner` has the same interface on sm_100a kernels.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch

from ..config import FedConfig, ModelConfig
from ..models.arena import FlatArena
from ..models.factory import build_model
from .adamw import FlatAdamW


def accuracy_from_logits(logits: torch.Tensor, labels: torch.Tensor) -> Tuple[int, int]:
    pred = torch.argmax(logits, dim=-1)
    return int((pred == labels).sum().item()), int(labels.numel())


class TorchRunner:
    """nn.Module + flat arena + FlatAdamW.  ``flat`` is the whole model as one fp32 vector."""

    backend = "torch"

    def __init__(self, model_cfg: ModelConfig, fed: Optional[FedConfig] = None, device="cpu", seed: Optional[int] = 0,
                 module: Optional[torch.nn.Module] = None):
        fed = fed or FedConfig()
        self.cfg, self.fed, self.device = model_cfg, fed, torch.device(device)
        if fed.deterministic:               # --deterministic on the PyTorch path: ask torch for its reproducible kernels
            torch.use_deterministic_algorithms(True, warn_only=True)
        self.module = module if module is not None else build_model(
            model_cfg, self.device, seed, checkpoint=fed.init_checkpoint, ignore_mismatched_sizes=fed.ignore_mismatched_sizes)
        self.arena = FlatArena.adopt(self.module, self.device)
        self.grads = self.arena.like()
        for (k, p) in self.module.named_parameters():
            p.grad = self.grads.views[k]
        self.opt = FlatAdamW(self.arena.flat, self.grads.flat, fed.lr, fed.adam_betas, fed.adam_eps,
                             fed.weight_decay, use_kernel=False)
        self.samples_seen = 0
        self.tokens_seen = 0

    # ---- parameters -------------------------------------------------------------------
    @property
    def flat(self) -> torch.Tensor:
        return self.arena.flat

    def state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        return self.arena.state_dict()

    def load_flat(self, flat: torch.Tensor) -> None:
        self.arena.flat.copy_(flat.to(self.arena.flat.device))

    def load_state_dict(self, state: Dict[str, torch.Tensor]) -> None:
        missing = set(self.arena.views) ^ set(state)
        if missing:
            raise KeyError(f"state_dict key mismatch (strict): {sorted(missing)[:4]} …")
        self.arena.load(state)

    def sync_shadow(self) -> None:   # interface parity with FusedRunner
        pass

    # ---- loops ------------------------------------------------------------------------
    def train_epochs(self, loader, epochs: int = 1, lazy: bool = False) -> Dict[str, float]:
        self.opt.reset()                      # AdamW(...) re-created each call (:109)
        self.module.train()
        last, n_steps = float("nan"), 0
        for _ in range(epochs):
            for batch in loader:
                batch = {k: v.to(self.device) for k, v in batch.items()}
                out = self.module(**batch)
                out.loss.backward()
                self.opt.step(zero_grad=True)
                last = float(out.loss.detach())
                n_steps += 1
                self.samples_seen += int(batch["labels"].numel())
                self.tokens_seen += int(batch["attention_mask"].sum())
        if lazy:                              # same contract as FusedRunner: a tensor the caller reads back later
            return {"loss": torch.tensor(last, dtype=torch.float32), "steps": n_steps}
        return {"loss": last, "steps": n_steps}

    def evaluate_lazy(self, loader):
        loss, acc = self.evaluate(loader)
        n_ds = max(len(loader.dataset), 1)
        total = sum(len(i) for i in loader.batch_indices())
        return (torch.tensor(loss * n_ds, dtype=torch.float32), torch.tensor(int(round(acc * total)), dtype=torch.int64),
                n_ds, total)

    @torch.no_grad()
    def evaluate(self, loader) -> Tuple[float, float]:
        self.module.eval()
        loss_sum, correct, total = 0.0, 0, 0
        for batch in loader:
            batch = {k: v.to(self.device) for k, v in batch.items()}
            out = self.module(**batch)
            loss_sum += float(out.loss)
            c, n = accuracy_from_logits(out.logits, batch["labels"])
            correct, total = correct + c, total + n
        n_ds = max(len(loader.dataset), 1)
        return loss_sum / n_ds, (correct / total if total else 0.0)


def resolve_device(fed: FedConfig) -> torch.device:
    if fed.device == "auto":
        return torch.device("cuda" if torch.cuda.is_available() else "cpu")
    return torch.device(fed.device)


def make_runner(fed: FedConfig, device=None, seed: Optional[int] = None):
    """Pick the fused sm_100a runner on a GPU (BERT family) or the torch runner otherwise."""
    dev = torch.device(device) if device is not None else resolve_device(fed)
    seed = fed.seed if seed is None else seed
    backend = fed.backend
    if backend == "auto":
        fusable = fed.model.arch in ("bert", "albert") and fed.model.head_dim == 64 and fed.model.hidden_size <= 1024
        backend = "fused" if (dev.type == "cuda" and fusable) else "torch"
    if backend == "fused":
        from ..models.fused_bert import FusedRunner
        return FusedRunner(fed.model, fed, dev, seed)
    return TorchRunner(fed.model, fed, dev, seed)
