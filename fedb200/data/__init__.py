"""Datasets, tokeniser, partitioners and the batch loader (reference layer L1, C1-C3, C27)."""
from .dataset import TokenDataset, BatchLoader, collate_padded  # noqa: F401
from .synthetic import make_synthetic, DATASET_SHAPES  # noqa: F401
from .partition import iid_indices, shard_indices, dirichlet_indices, Partitioner  # noqa: F401
from .tokenizer import HashTokenizer  # noqa: F401
from .csv_datasets import load_csv_dataset  # noqa: F401
from .build import build_corpus, client_loaders  # noqa: F401
