"""Tokenised dataset container + batch loader with HF ``DataCollatorWithPadding`` semantics.

Reference pipeline (src/Servercase/server_IID_IMDB.py:65-101): tokenise with truncation,
``select`` a subset, drop text columns, rename label->``labels``, pad each batch to its
longest member, batch 32, shuffle only the train loader.  ``len(loader)`` is the number
of *batches* — the reference uses that as the FedAvg weight (server_IID_IMDB.py:174).
"""
from __future__ import annotations

from typing import Dict, Iterator, List, Optional, Sequence

import numpy as np
import torch


class TokenDataset:
    """Variable-length token sequences stored packed: ``tokens[offsets[i]:offsets[i+1]]``."""

    def __init__(self, tokens: np.ndarray, offsets: np.ndarray, labels: np.ndarray, num_labels: int,
                 pad_id: int = 0):
        self.tokens = np.ascontiguousarray(tokens, dtype=np.int32)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        self.labels = np.ascontiguousarray(labels, dtype=np.int64)
        self.num_labels = int(num_labels)
        self.pad_id = pad_id
        assert len(self.offsets) == len(self.labels) + 1

    @classmethod
    def from_sequences(cls, seqs: Sequence[Sequence[int]], labels: Sequence[int], num_labels: int, pad_id=0):
        lens = np.fromiter((len(s) for s in seqs), dtype=np.int64, count=len(seqs))
        offsets = np.zeros(len(seqs) + 1, dtype=np.int64)
        np.cumsum(lens, out=offsets[1:])
        tokens = np.concatenate([np.asarray(s, dtype=np.int32) for s in seqs]) if len(seqs) else np.zeros(0, np.int32)
        return cls(tokens, offsets, np.asarray(labels), num_labels, pad_id)

    def __len__(self) -> int:
        return len(self.labels)

    def lengths(self) -> np.ndarray:
        return np.diff(self.offsets)

    def sequence(self, i: int) -> np.ndarray:
        return self.tokens[self.offsets[i]:self.offsets[i + 1]]

    def select(self, indices: Sequence[int]) -> "TokenDataset":
        """HF ``Dataset.select`` equivalent (server_IID_IMDB.py:83-84)."""
        idx = np.asarray(indices, dtype=np.int64)
        seqs = [self.sequence(int(i)) for i in idx]
        return TokenDataset.from_sequences(seqs, self.labels[idx], self.num_labels, self.pad_id)


def collate_padded(ds: TokenDataset, idx: np.ndarray, pin: bool = False) -> Dict[str, torch.Tensor]:
    """Pad to the longest member of the batch (``DataCollatorWithPadding``, :89)."""
    lens = (ds.offsets[idx + 1] - ds.offsets[idx]).astype(np.int64)
    B, S = len(idx), int(lens.max()) if len(idx) else 0
    ids = np.full((B, S), ds.pad_id, dtype=np.int64)
    mask = np.zeros((B, S), dtype=np.int64)
    for r, i in enumerate(idx):
        n = lens[r]
        ids[r, :n] = ds.tokens[ds.offsets[i]:ds.offsets[i] + n]
        mask[r, :n] = 1
    out = {
        "input_ids": torch.from_numpy(ids),
        "token_type_ids": torch.zeros((B, S), dtype=torch.int64),
        "attention_mask": torch.from_numpy(mask),
        "labels": torch.from_numpy(ds.labels[idx].copy()),
    }
    if pin and torch.cuda.is_available():
        out = {k: v.pin_memory() for k, v in out.items()}
    return out


def collate_packed(ds: TokenDataset, idx: np.ndarray, pin: bool = False, longest_first: bool = False) -> Dict[str, torch.Tensor]:
    """Packed (pad-free) collation for the fused trainer: ``ids/pos/tt`` are ``[T]`` int32 with the
    sequences laid end to end, ``cu_seqlens`` is ``[B+1]``.  Same samples as :func:`collate_padded`, none
    of the pad tokens.  ``longest_first`` lays the sequences out by decreasing length: the attention kernels launch one
    CTA per (tile, head, sequence) in sequence order, so the long sequences' CTAs start first and the short ones fill the
    tail of the grid (a training batch's loss is a mean over sequences, so their order is free)."""
    idx = np.asarray(idx)
    perm = None
    if longest_first and len(idx) > 1:
        perm = np.argsort(-(ds.offsets[idx + 1] - ds.offsets[idx]), kind="stable")     # packed position -> caller's position
        idx = idx[perm]
    lens = (ds.offsets[idx + 1] - ds.offsets[idx]).astype(np.int64)
    B, T = len(idx), int(lens.sum())
    cu = np.zeros(B + 1, dtype=np.int32)
    np.cumsum(lens, out=cu[1:])
    ids = np.empty(T, dtype=np.int32)
    pos = np.empty(T, dtype=np.int32)
    for r, i in enumerate(idx):
        n = int(lens[r])
        ids[cu[r]:cu[r] + n] = ds.tokens[ds.offsets[i]:ds.offsets[i] + n]
        pos[cu[r]:cu[r] + n] = np.arange(n, dtype=np.int32)
    out = {"ids": torch.from_numpy(ids), "pos": torch.from_numpy(pos), "tt": torch.zeros(T, dtype=torch.int32),
           "cu_seqlens": torch.from_numpy(cu), "labels": torch.from_numpy(ds.labels[idx].astype(np.int32))}
    if pin and torch.cuda.is_available():
        out = {k: v.pin_memory() for k, v in out.items()}
    if perm is not None:
        pt = torch.from_numpy(perm.astype(np.int64))
        out["perm"] = pt.pin_memory() if (pin and torch.cuda.is_available()) else pt
    out["max_len"] = int(lens.max()) if B else 0
    out["num_tokens"] = T
    return out


class BatchLoader:
    """Minimal ``DataLoader`` replacement: batch 32, optional shuffle, padded collation.

    ``dataset`` attribute mirrors ``torch.utils.data.DataLoader.dataset`` because the
    reference's evaluator divides by ``len(testloader.dataset)`` (server_IID_IMDB.py:133).
    """

    def __init__(self, dataset: TokenDataset, batch_size: int = 32, shuffle: bool = False,
                 rng: Optional[np.random.Generator] = None, pin: bool = False):
        self.dataset = dataset
        self.batch_size = int(batch_size)
        self.shuffle = shuffle
        self.rng = rng or np.random.default_rng()
        self.pin = pin

    def __len__(self) -> int:
        return (len(self.dataset) + self.batch_size - 1) // self.batch_size

    def batch_indices(self) -> List[np.ndarray]:
        n = len(self.dataset)
        order = self.rng.permutation(n) if self.shuffle else np.arange(n)
        return [order[i:i + self.batch_size] for i in range(0, n, self.batch_size)]

    def __iter__(self) -> Iterator[Dict[str, torch.Tensor]]:
        for idx in self.batch_indices():
            yield collate_padded(self.dataset, idx, pin=self.pin)

    def packed(self) -> Iterator[Dict[str, torch.Tensor]]:
        for idx in self.batch_indices():
            yield collate_packed(self.dataset, idx, pin=self.pin)

    @property
    def num_samples(self) -> int:
        return len(self.dataset)

    def num_tokens(self) -> int:
        return int(self.dataset.lengths().sum())
