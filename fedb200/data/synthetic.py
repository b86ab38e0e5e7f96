"""Synthetic corpora with the *shape* of the reference's datasets (no network, no hub).

Shapes: IMDB = 2 classes, 25 000 train / 25 000 test, reviews mostly 150-512 word pieces
(truncated at 512, src/Servercase/server_IID_IMDB.py:67-76); medical transcriptions =
40 classes, 12 000 / 3 000 rows, descriptions ~25-100 word pieces (Dataset/train_file_mt.csv);
cancer / covid = 41 labels (src/Serverlesscase/serverless_caner_classification_iid.py:120);
self-driving sentiment = 3 labels, 500 rows (Dataset/sentiment_analysis_self_driving_vehicles.csv).

Documents are class-conditional bags of tokens so a classifier can actually learn them:
each class owns a small set of "signal" tokens drawn with elevated probability.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .dataset import TokenDataset


@dataclass(frozen=True)
class DatasetShape:
    num_labels: int
    train_rows: int
    test_rows: int
    len_mean: float       # mean of log-normal body length (tokens)
    len_sigma: float
    len_min: int
    len_max: int
    label_sorted: bool    # HF imdb train split is label-sorted (serverless_NonIID_IMDB.py:48 relies on it)


DATASET_SHAPES = {
    "imdb": DatasetShape(2, 25000, 25000, 300.0, 0.55, 32, 510, True),
    "medical_transcriptions": DatasetShape(40, 12000, 3000, 45.0, 0.45, 8, 100, False),
    "cancer": DatasetShape(41, 5408, 1352, 220.0, 0.5, 32, 510, False),
    "covid": DatasetShape(41, 8000, 2000, 60.0, 0.5, 8, 200, False),
    "self_driving": DatasetShape(3, 500, 500, 24.0, 0.3, 6, 64, False),
}


def make_synthetic(name: str, split: str = "train", vocab_size: int = 28996, max_seq_len: int = 512,
                   rows: int | None = None, seed: int = 42, signal_tokens: int = 24,
                   signal_rate: float = 0.12, shuffled: bool = True) -> TokenDataset:
    """Build a synthetic ``TokenDataset`` for ``name``.

    ``shuffled=False`` keeps the natural order (label-sorted for IMDB) which the contiguous
    shard partitioner depends on for its pathological label skew.
    """
    shp = DATASET_SHAPES[name]
    n = rows if rows is not None else (shp.train_rows if split == "train" else shp.test_rows)
    rng = np.random.default_rng(seed + (0 if split == "train" else 7919))
    if shp.label_sorted:
        labels = (np.arange(n) * shp.num_labels // max(n, 1)).astype(np.int64)
    else:
        labels = rng.integers(0, shp.num_labels, size=n)
    body_max = min(shp.len_max, max_seq_len - 2)
    lens = np.exp(rng.normal(np.log(shp.len_mean), shp.len_sigma, size=n))
    lens = np.clip(lens, shp.len_min, body_max).astype(np.int64)
    # class signal vocab: fixed per dataset (independent of split seed)
    srng = np.random.default_rng(1234567 + shp.num_labels)
    lo = 1000 if vocab_size > 2000 else 8
    sig = srng.integers(lo, vocab_size, size=(shp.num_labels, signal_tokens))
    offsets = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens + 2, out=offsets[1:])
    tokens = rng.integers(lo, vocab_size, size=int(offsets[-1]), dtype=np.int64).astype(np.int32)
    is_sig = rng.random(int(offsets[-1])) < signal_rate
    pick = rng.integers(0, signal_tokens, size=int(offsets[-1]))
    row_of = np.repeat(np.arange(n), lens + 2)
    tokens = np.where(is_sig, sig[labels[row_of], pick], tokens).astype(np.int32)
    cls_id, sep_id = (101, 102) if vocab_size > 2000 else (1, 2)
    tokens[offsets[:-1]] = cls_id
    tokens[offsets[1:] - 1] = sep_id
    ds = TokenDataset(tokens, offsets, labels, shp.num_labels)
    if shuffled:
        perm = np.random.default_rng(seed).permutation(n)   # raw_datasets.shuffle(seed=42)
        ds = ds.select(perm)
    return ds
