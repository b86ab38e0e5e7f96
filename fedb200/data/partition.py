"""Client partitioners (C2, C3 + the Dirichlet extension named by BASELINE.json).

* IID: a fresh random subset per client visit — ``random.sample(range(len(split)), n)``
  (src/Servercase/server_IID_IMDB.py:79-84; resampled every round in
  src/Serverlesscase/serverless_IID_IMDB.py:258).
* Contiguous shard ("non-IID"): client *i* trains on rows ``[stride*i, stride*i+train_n)`` of
  the **unshuffled** split; IMDB uses stride 300 / 240 train with the test slice being the
  tail of the same stride (src/Serverlesscase/serverless_NonIID_IMDB.py:59-60); medical uses
  stride 500 / 400 train and fixed test rows 0..400
  (src/Serverlesscase/Serverless_NonIID_Medical_transcriptions.py:55-56).
* Dirichlet label skew: not in the reference; per-class proportions ~ Dir(alpha).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np


def iid_indices(population: int, n: int, rng: np.random.Generator) -> np.ndarray:
    """Sample ``n`` distinct row ids without replacement (``random.sample`` semantics)."""
    n = min(n, population)
    return rng.choice(population, size=n, replace=False)


def shard_indices(client: int, stride: int = 300, train_n: int = 240, test_mode: str = "tail",
                  test_n: Optional[int] = None) -> Tuple[np.ndarray, np.ndarray]:
    """Contiguous shard index math of the reference's non-IID scripts.

    ``test_mode='tail'``: test rows ``[stride*i+train_n, stride*(i+1))`` (IMDB variant);
    ``test_mode='fixed'``: test rows ``[0, test_n)`` for every client (medical variant).
    """
    lo = stride * client
    train = np.arange(lo, lo + train_n)
    if test_mode == "tail":
        test = np.arange(lo + train_n, lo + stride)
    elif test_mode == "fixed":
        test = np.arange(0, test_n if test_n is not None else train_n)
    else:
        raise ValueError(test_mode)
    return train, test


def dirichlet_indices(labels: np.ndarray, num_clients: int, alpha: float, rng: np.random.Generator,
                      min_per_client: int = 1) -> List[np.ndarray]:
    """Label-skewed split: for each class, divide its rows among clients by Dir(alpha)."""
    labels = np.asarray(labels)
    classes = np.unique(labels)
    for _ in range(100):
        parts: List[List[int]] = [[] for _ in range(num_clients)]
        for c in classes:
            rows = np.flatnonzero(labels == c)
            rng.shuffle(rows)
            p = rng.dirichlet(np.full(num_clients, alpha))
            cuts = (np.cumsum(p) * len(rows)).astype(int)[:-1]
            for k, chunk in enumerate(np.split(rows, cuts)):
                parts[k].extend(chunk.tolist())
        if min(len(p) for p in parts) >= min_per_client:
            break
    return [np.asarray(sorted(p), dtype=np.int64) for p in parts]


@dataclass
class Partitioner:
    """Maps (client, round) -> (train row ids, test row ids) into the corpus."""
    kind: str                    # iid | shard | dirichlet
    train_population: int
    test_population: int
    train_n: int = 100
    test_n: int = 100
    stride: int = 300
    shard_train: int = 240
    shard_test_mode: str = "tail"
    seed: int = 0
    _dirichlet: Optional[List[np.ndarray]] = None

    def prepare_dirichlet(self, labels: np.ndarray, num_clients: int, alpha: float) -> None:
        self._dirichlet = dirichlet_indices(labels, num_clients, alpha, np.random.default_rng(self.seed))

    def indices(self, client: int, round_idx: int) -> Tuple[np.ndarray, np.ndarray]:
        rng = np.random.default_rng([self.seed, client, round_idx, 0x5EED])
        if self.kind == "iid":
            return (iid_indices(self.train_population, self.train_n, rng),
                    iid_indices(self.test_population, self.test_n, rng))
        if self.kind == "shard":
            tr, te = shard_indices(client, self.stride, self.shard_train, self.shard_test_mode, self.test_n)
            return tr % self.train_population, te % (self.train_population if self.shard_test_mode == "tail"
                                                     else self.test_population)
        if self.kind == "dirichlet":
            assert self._dirichlet is not None, "call prepare_dirichlet first"
            pool = self._dirichlet[client % len(self._dirichlet)]
            take = min(self.train_n, len(pool))
            return rng.choice(pool, size=take, replace=False), iid_indices(self.test_population, self.test_n, rng)
        raise ValueError(self.kind)
