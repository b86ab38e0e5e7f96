"""fedb200 — Blackwell-native federated fine-tuning for BERT-family classifiers.

Server-mode FedAvg and serverless peer-to-peer gossip, IID / non-IID client splits,
synchronous and asynchronous rounds, an optimised weight-transfer route, PageRank
elimination of anomalous peers and a hash-chained ledger of every weight update.

The compute path is PyTorch tensors + hand-written sm_100a kernels (``csrc/``, loaded
through :mod:`fedb200.ops._lib`); everything else here is the host runtime around it.
Capability source: the research scripts under ``src/Servercase`` and
``src/Serverlesscase`` of the reference repository (see SURVEY.md for file:line).
"""
__version__ = "0.1.0"

from .config import FedConfig, ModelConfig, DataConfig  # noqa: F401
