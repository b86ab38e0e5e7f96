"""One dataclass configuration + CLI shared by all entrypoints.

The reference has no flag system: it edits module constants ``DEVICE / CHECKPOINT /
NUM_CLIENTS / NUM_ROUNDS`` (src/Servercase/server_IID_IMDB.py:47-50) and buries the rest
as literals (batch 32 :93, lr 5e-5 :109, 1 epoch :172, 100 samples :79-80).  Those
literals are the defaults here.
"""
from __future__ import annotations

import argparse
import dataclasses
import json
from dataclasses import dataclass, field
from typing import Optional

# Architecture presets.  ``biobert`` is dmis-lab/biobert-v1.1 (BERT-base-cased shape,
# vocab 28 996 — serverless_cancer_classification_with_BioBERT.ipynb:526-570);
# ``albert`` is albert-base-v2 (src/Serverlesscase/serverless_IID_IMDB.py:31).
MODEL_PRESETS = {
    "bert-base": dict(arch="bert", vocab_size=30522, hidden_size=768, num_layers=12, num_heads=12,
                      intermediate_size=3072),
    "biobert": dict(arch="bert", vocab_size=28996, hidden_size=768, num_layers=12, num_heads=12,
                    intermediate_size=3072),
    "bert-large": dict(arch="bert", vocab_size=30522, hidden_size=1024, num_layers=24, num_heads=16,
                       intermediate_size=4096),
    "albert-base-v2": dict(arch="albert", vocab_size=30000, hidden_size=768, num_layers=12, num_heads=12,
                           intermediate_size=3072, embedding_size=128),
    # small head_dim-64 shape for GPU tests of the fused trainer
    "bert-mini": dict(arch="bert", vocab_size=2048, hidden_size=128, num_layers=2, num_heads=2,
                      intermediate_size=256, max_position=512),
    "albert-mini": dict(arch="albert", vocab_size=2048, hidden_size=128, num_layers=3, num_heads=2,
                        intermediate_size=256, embedding_size=64, max_position=512),
    # tiny shapes for CPU tests
    "bert-tiny": dict(arch="bert", vocab_size=512, hidden_size=64, num_layers=2, num_heads=2,
                      intermediate_size=128, max_position=128),
    "albert-tiny": dict(arch="albert", vocab_size=512, hidden_size=64, num_layers=2, num_heads=2,
                        intermediate_size=128, embedding_size=32, max_position=128),
}


@dataclass
class ModelConfig:
    arch: str = "bert"
    vocab_size: int = 28996
    hidden_size: int = 768
    num_layers: int = 12
    num_heads: int = 12
    intermediate_size: int = 3072
    max_position: int = 512
    type_vocab_size: int = 2
    embedding_size: int = 128          # ALBERT factorised embedding
    num_labels: int = 2
    layer_norm_eps: float = 1e-12
    hidden_dropout: float = 0.1
    attention_dropout: float = 0.1
    initializer_range: float = 0.02
    name: str = "biobert"

    @classmethod
    def preset(cls, name: str, **overrides) -> "ModelConfig":
        if name not in MODEL_PRESETS:
            raise KeyError(f"unknown model preset {name!r}; have {sorted(MODEL_PRESETS)}")
        kw = dict(MODEL_PRESETS[name])
        kw.update(overrides)
        return cls(name=name, **kw)

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_heads


@dataclass
class DataConfig:
    dataset: str = "imdb"              # imdb | medical_transcriptions | cancer | covid | self_driving
    source: str = "synthetic"          # synthetic | csv
    data_dir: Optional[str] = None     # directory with the bundled CSVs (Dataset/)
    split: str = "iid"                 # iid | shard | dirichlet
    train_per_client: int = 100        # server_IID_IMDB.py:79
    test_per_client: int = 100         # server_IID_IMDB.py:80
    batch_size: int = 32               # server_IID_IMDB.py:93
    max_seq_len: int = 512             # tokenizer truncation=True => BERT max 512
    shard_stride: int = 300            # serverless_NonIID_IMDB.py:59-60
    shard_train: int = 240
    dirichlet_alpha: float = 0.3
    population: int = 25000            # size of the synthetic corpus the partitioners index into
    seed: int = 42                     # raw_datasets.shuffle(seed=42), server_IID_IMDB.py:68


@dataclass
class FedConfig:
    mode: str = "serverless"           # server | serverless
    num_clients: int = 10
    num_rounds: int = 20
    local_epochs: int = 1              # server_IID_IMDB.py:172
    lr: float = 5e-5                   # server_IID_IMDB.py:109
    adam_eps: float = 1e-6             # transformers.AdamW default (reference import, :36)
    adam_betas: tuple = (0.9, 0.999)
    weight_decay: float = 0.0
    device: str = "auto"               # auto | cpu | cuda
    dtype: str = "bf16"                # compute dtype of the fused trainer (master weights fp32)
    backend: str = "auto"              # auto | fused (sm_100a kernels) | torch (plain modules)
    comm: str = "auto"                 # auto | loopback | peer | nccl
    # serverless semantics: "parallel" = every peer starts from the common global model;
    # "handoff" = the reference's sequential chain (serverless_IID_IMDB.py:255-266).
    serverless_semantics: str = "parallel"
    asynchronous: bool = False
    staleness_decay: float = 0.5       # alpha(tau) = base / (1 + tau)^decay
    async_scheduled: bool = False      # force the deterministic schedule-driven executor (always used when world > 1)
    async_epoch_s: float = 0.05        # modelled local-epoch duration of the schedule's virtual clock
    async_spread: float = 0.5          # peer heterogeneity of the modelled durations
    async_fanout: Optional[int] = None # destinations per publish (None = every reachable peer)
    # multi-GPU asynchronous executor: "observed" = every mix takes whatever versions its sources have published by then
    # (device-side decision, chunk flags, no schedule, nobody waits for a peer); "scheduled" = the deterministic virtual-time
    # schedule of round 1 (modelled epoch durations)
    async_mode: str = "auto"           # auto: observed across GPUs, the virtual-time event engine in one process
    async_fanin: int = 4               # sources per mix in observed mode (bounded fan-in: a mix reads <= 4 remote models)
    async_ring: int = 3                # outbox versions kept per peer (slot = version % ring)
    async_chunk_kb: int = 4096         # landing-flag granularity of a publish (KB)
    async_score_every: int = 0         # observed mode + cosine_pagerank: re-score every K epochs (0 = never)
    async_time: str = "measured"       # single-process engine clock: measured epoch durations | modelled (async_epoch_s, async_spread)
    async_min_latency_s: float = 0.0   # visibility lag of a publish in the schedule's virtual time (see AsyncGossipEngine)
    topology: str = "full"             # full | ring | reference10 | random
    route_optimizer: bool = True
    anomaly_detector: str = "none"     # none | pagerank | modz | dbscan | community | cosine_pagerank
    inject_anomalies: int = 0
    inject_kind: str = "scale"         # scale | signflip | noise
    ledger: bool = False
    verify_updates: bool = False       # receivers recompute the Merkle root of an update they consume and compare it with the recorded one
    ledger_chunk_bytes: int = 1 << 20
    # fine-tune from a local HF checkpoint directory (config.json + model.safetensors / pytorch_model.bin [+ vocab.txt]):
    # AutoModelForSequenceClassification.from_pretrained(CHECKPOINT, num_labels=...), server_IID_IMDB.py:142-144
    init_checkpoint: Optional[str] = None
    ignore_mismatched_sizes: bool = False   # server_noniid_medical_transcriptions.py:146-148
    vocab_path: Optional[str] = None   # vocab.txt (or its directory) for the WordPiece tokenizer; default: init_checkpoint
    deterministic: bool = False        # bit-reproducible training (csrc/determinism.cu: ordered reductions instead of fp32 atomics;
                                       # +~10 % step time): the ledger's Merkle roots can be re-derived by re-running a round
    nvls: bool = False                 # FedAvg reduce inside the NVSwitch (multimem) where the box exposes multicast
    aggregate_ctas: int = 32           # grid of the fused FedAvg reduce when it overlaps the local evaluation (side stream)
    placement: str = "balanced"        # balanced (token-count LPT) | block (contiguous id blocks)
    checkpoint_dir: Optional[str] = None
    checkpoint_async: bool = True      # background writer: the round loop only pays a device->pinned-host copy
    checkpoint_every: int = 1          # serverless_IID_IMDB.py:280 saves every round
    resume: bool = False
    metrics_path: Optional[str] = None
    seed: int = 0
    model: ModelConfig = field(default_factory=ModelConfig)
    data: DataConfig = field(default_factory=DataConfig)

    def to_json(self) -> str:
        return json.dumps(dataclasses.asdict(self), indent=1, default=str)


def build_parser(defaults: Optional[FedConfig] = None, description: str = "") -> argparse.ArgumentParser:
    d = defaults or FedConfig()
    p = argparse.ArgumentParser(description=description)
    p.add_argument("--model", default=d.model.name, choices=sorted(MODEL_PRESETS))
    p.add_argument("--num-labels", type=int, default=None)
    p.add_argument("--dataset", default=d.data.dataset)
    p.add_argument("--source", default=d.data.source, choices=["synthetic", "csv"])
    p.add_argument("--data-dir", default=d.data.data_dir)
    p.add_argument("--split", default=d.data.split, choices=["iid", "shard", "dirichlet"])
    p.add_argument("--dirichlet-alpha", type=float, default=d.data.dirichlet_alpha)
    p.add_argument("--clients", type=int, default=d.num_clients)
    p.add_argument("--rounds", type=int, default=d.num_rounds)
    p.add_argument("--train-per-client", type=int, default=d.data.train_per_client)
    p.add_argument("--test-per-client", type=int, default=d.data.test_per_client)
    p.add_argument("--batch-size", type=int, default=d.data.batch_size)
    p.add_argument("--max-seq-len", type=int, default=d.data.max_seq_len)
    p.add_argument("--lr", type=float, default=d.lr)
    p.add_argument("--device", default=d.device)
    p.add_argument("--backend", default=d.backend, choices=["auto", "fused", "torch"])
    p.add_argument("--comm", default=d.comm, choices=["auto", "loopback", "peer", "nccl"])
    p.add_argument("--semantics", default=d.serverless_semantics, choices=["parallel", "handoff"])
    p.add_argument("--async", dest="asynchronous", action="store_true", default=d.asynchronous)
    p.add_argument("--async-scheduled", action="store_true", default=d.async_scheduled)
    p.add_argument("--async-fanout", type=int, default=d.async_fanout)
    p.add_argument("--async-mode", default=d.async_mode, choices=["auto", "observed", "scheduled", "engine"])
    p.add_argument("--async-fanin", type=int, default=d.async_fanin)
    p.add_argument("--async-ring", type=int, default=d.async_ring)
    p.add_argument("--async-score-every", type=int, default=d.async_score_every)
    p.add_argument("--async-epoch-s", type=float, default=d.async_epoch_s)
    p.add_argument("--async-min-latency-s", type=float, default=d.async_min_latency_s)
    p.add_argument("--async-time", default=d.async_time, choices=["measured", "modelled"])
    p.add_argument("--topology", default=d.topology)
    p.add_argument("--no-route-optimizer", dest="route_optimizer", action="store_false", default=d.route_optimizer)
    p.add_argument("--anomaly-detector", default=d.anomaly_detector)
    p.add_argument("--inject-anomalies", type=int, default=d.inject_anomalies)
    p.add_argument("--inject-kind", default=d.inject_kind)
    p.add_argument("--ledger", action="store_true", default=d.ledger)
    p.add_argument("--verify-updates", action="store_true", default=d.verify_updates)
    p.add_argument("--nvls", action="store_true", default=d.nvls, help="switch-side FedAvg reduction (NVLS multicast)")
    p.add_argument("--init-checkpoint", default=d.init_checkpoint,
                   help="local HF checkpoint directory to fine-tune from (from_pretrained equivalent)")
    p.add_argument("--ignore-mismatched-sizes", action="store_true", default=d.ignore_mismatched_sizes)
    p.add_argument("--vocab", dest="vocab_path", default=d.vocab_path, help="vocab.txt for the WordPiece tokenizer")
    p.add_argument("--deterministic", action="store_true", default=d.deterministic)
    p.add_argument("--placement", default=d.placement, choices=["balanced", "block"])
    p.add_argument("--checkpoint-sync", dest="checkpoint_async", action="store_false", default=d.checkpoint_async)
    p.add_argument("--checkpoint-dir", default=d.checkpoint_dir)
    p.add_argument("--checkpoint-every", type=int, default=d.checkpoint_every)
    p.add_argument("--resume", action="store_true", default=d.resume)
    p.add_argument("--metrics", dest="metrics_path", default=d.metrics_path)
    p.add_argument("--seed", type=int, default=d.seed)
    p.add_argument("--no-dropout", action="store_true")
    return p


DATASET_LABELS = {"imdb": 2, "medical_transcriptions": 40, "cancer": 41, "covid": 41, "self_driving": 3}


def config_from_args(args: argparse.Namespace, defaults: Optional[FedConfig] = None) -> FedConfig:
    cfg = dataclasses.replace(defaults) if defaults else FedConfig()
    nl = args.num_labels if args.num_labels is not None else DATASET_LABELS.get(args.dataset, 2)
    mo = dict(num_labels=nl)
    if args.no_dropout:
        mo.update(hidden_dropout=0.0, attention_dropout=0.0)
    cfg.model = ModelConfig.preset(args.model, **mo)
    cfg.init_checkpoint, cfg.ignore_mismatched_sizes = args.init_checkpoint, args.ignore_mismatched_sizes
    cfg.vocab_path = args.vocab_path or args.init_checkpoint
    cfg.deterministic, cfg.placement, cfg.checkpoint_async = args.deterministic, args.placement, args.checkpoint_async
    if args.init_checkpoint:
        # the architecture comes from the checkpoint's config.json, the label count from the task (num_labels=...)
        import dataclasses as _dc
        from .models.hf_io import load_config
        loaded = load_config(args.init_checkpoint)
        drop = dict(hidden_dropout=0.0, attention_dropout=0.0) if args.no_dropout else {}
        cfg.model = _dc.replace(loaded, num_labels=nl, name=args.model, **drop)
    dd = dataclasses.asdict(cfg.data)
    dd.update(dataset=args.dataset, source=args.source, data_dir=args.data_dir, split=args.split,
              dirichlet_alpha=args.dirichlet_alpha, train_per_client=args.train_per_client,
              test_per_client=args.test_per_client, batch_size=args.batch_size,
              max_seq_len=min(args.max_seq_len, cfg.model.max_position))
    cfg.data = DataConfig(**dd)
    cfg.num_clients, cfg.num_rounds, cfg.lr = args.clients, args.rounds, args.lr
    cfg.device, cfg.backend, cfg.comm = args.device, args.backend, args.comm
    cfg.serverless_semantics, cfg.asynchronous = args.semantics, args.asynchronous
    cfg.async_scheduled, cfg.async_fanout, cfg.async_epoch_s = args.async_scheduled, args.async_fanout, args.async_epoch_s
    cfg.async_min_latency_s, cfg.async_time = args.async_min_latency_s, args.async_time
    cfg.async_mode, cfg.async_fanin, cfg.async_ring = args.async_mode, args.async_fanin, args.async_ring
    cfg.async_score_every = args.async_score_every
    cfg.topology, cfg.route_optimizer = args.topology, args.route_optimizer
    cfg.anomaly_detector, cfg.inject_anomalies, cfg.inject_kind = (
        args.anomaly_detector, args.inject_anomalies, args.inject_kind)
    cfg.ledger, cfg.checkpoint_dir, cfg.checkpoint_every = args.ledger, args.checkpoint_dir, args.checkpoint_every
    cfg.verify_updates, cfg.nvls = args.verify_updates, args.nvls
    cfg.resume, cfg.metrics_path, cfg.seed = args.resume, args.metrics_path, args.seed
    return cfg
