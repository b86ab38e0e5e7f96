"""Corpus + per-client loader construction used by every entrypoint.

Equivalent of the reference's ``load_data`` family (src/Servercase/server_IID_IMDB.py:65-101,
``load_data_clients(i)`` src/Serverlesscase/serverless_NonIID_IMDB.py:45-77) but the
corpus is tokenised once and only indexed per client visit — the reference re-downloads and
re-tokenises the whole dataset on every visit.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple

import numpy as np

from ..config import DataConfig, ModelConfig
from .csv_datasets import CSV_SPECS, load_csv_dataset
from .dataset import BatchLoader, TokenDataset
from .partition import Partitioner
from .synthetic import DATASET_SHAPES, make_synthetic


@dataclass
class Corpus:
    train: TokenDataset
    test: TokenDataset
    train_unshuffled: TokenDataset
    partitioner: Partitioner


def build_corpus(data: DataConfig, model: ModelConfig, num_clients: int, seed: int = 0,
                 vocab_path: str | None = None) -> Corpus:
    if data.source == "csv":
        if not data.data_dir or data.dataset not in CSV_SPECS:
            raise ValueError("csv source needs --data-dir and one of " + ", ".join(CSV_SPECS))
        raw_train = load_csv_dataset(data.dataset, data.data_dir, "train", model.vocab_size, data.max_seq_len,
                                     vocab_path=vocab_path)
        raw_test = load_csv_dataset(data.dataset, data.data_dir, "test", model.vocab_size, data.max_seq_len,
                                    vocab_path=vocab_path)
        perm = np.random.default_rng(data.seed).permutation(len(raw_train))
        train, unshuf = raw_train.select(perm), raw_train
        test = raw_test
    else:
        shp = DATASET_SHAPES[data.dataset]
        rows_tr = min(shp.train_rows, data.population)
        rows_te = min(shp.test_rows, data.population)
        unshuf = make_synthetic(data.dataset, "train", model.vocab_size, data.max_seq_len, rows_tr, data.seed,
                                shuffled=False)
        perm = np.random.default_rng(data.seed).permutation(len(unshuf))
        train = unshuf.select(perm)
        test = make_synthetic(data.dataset, "test", model.vocab_size, data.max_seq_len, rows_te, data.seed)
    medical = data.dataset == "medical_transcriptions"
    part = Partitioner(kind=data.split, train_population=len(train), test_population=len(test),
                       train_n=data.train_per_client, test_n=data.test_per_client,
                       stride=500 if medical and data.shard_stride == 300 else data.shard_stride,
                       shard_train=400 if medical and data.shard_train == 240 else data.shard_train,
                       shard_test_mode="fixed" if medical else "tail", seed=seed)
    if data.split == "dirichlet":
        part.prepare_dirichlet(train.labels, num_clients, data.dirichlet_alpha)
    return Corpus(train, test, unshuf, part)


def client_loaders(corpus: Corpus, data: DataConfig, client: int, round_idx: int,
                   pin: bool = False) -> Tuple[BatchLoader, BatchLoader]:
    """(trainloader, testloader) for one client visit; train is shuffled, test is not (:90-99)."""
    tr_idx, te_idx = corpus.partitioner.indices(client, round_idx)
    if data.split == "shard":
        src = corpus.train_unshuffled            # non-IID scripts skip .shuffle (serverless_NonIID_IMDB.py:48)
        tr = src.select(tr_idx)
        te = (src if corpus.partitioner.shard_test_mode == "tail" else corpus.test).select(te_idx)
    else:
        tr, te = corpus.train.select(tr_idx), corpus.test.select(te_idx)
    rng = np.random.default_rng([corpus.partitioner.seed, client, round_idx, 0xBA7C])
    return (BatchLoader(tr, data.batch_size, shuffle=True, rng=rng, pin=pin),
            BatchLoader(te, data.batch_size, shuffle=False, pin=pin))


def global_test_loader(corpus: Corpus, data: DataConfig, round_idx: int, pin: bool = False) -> BatchLoader:
    """Fresh random test subset for the global evaluation (serverless_IID_IMDB.py:276-279)."""
    rng = np.random.default_rng([corpus.partitioner.seed, 0xFFFF, round_idx])
    idx = rng.choice(len(corpus.test), size=min(data.test_per_client, len(corpus.test)), replace=False)
    return BatchLoader(corpus.test.select(idx), data.batch_size, shuffle=False, pin=pin)
