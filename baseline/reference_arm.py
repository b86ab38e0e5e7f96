"""Reference arm of ``bench.py`` (``--impl reference``): runs the reference's OWN serverless script.

What is executed
----------------
``baseline/_ref/src/Serverlesscase/serverless_IID_IMDB.py`` — a byte-for-byte copy of
``/root/reference/src/Serverlesscase/serverless_IID_IMDB.py`` (checked against ``baseline/_ref/SHA256SUMS`` at
run time; staged by ``baseline/install_reference.sh`` because ``pip install`` of the reference is impossible:
it has no ``setup.py`` / ``pyproject.toml``).  The script is a flat module whose round loop is top-level code
(:255-285), so "calling its public API" means executing the module.  Every statement of it runs as written:
``load_data`` (:45-82), ``IMDBClient.train_model`` / ``get_parameters`` / ``evaluate_model`` (:155-187), the
host-numpy averaging and ``load_state_dict`` (:269-270), the global evaluation (:276-279) and
``save_pretrained`` (:280).  None of this repository's models, kernels or engine is on that path.

What had to be supplied for it to run offline (and why it does not change what is measured)
-------------------------------------------------------------------------------------------
* four module constants are re-bound when the source is compiled (``DEVICE``, ``CHECKPOINT``, ``NUM_CLIENTS``,
  ``NUM_ROUNDS`` — the reference's only "flags"), to the bench configuration: cuda, BioBERT shape,
  N peers, W + K rounds.  Nothing else in the syntax tree is touched.
* network / missing dependencies get offline stand-ins, installed into ``sys.modules`` before the script runs:
  - ``datasets.load_dataset("imdb")`` -> the synthetic IMDB-shape corpus (``fedb200.data.synthetic``: the same rows
    the fedb200 arm trains on) rendered as text; ``.map(tokenize)`` is deferred to the rows ``.select`` keeps
    (what the ``datasets`` on-disk cache does for the reference after its first call) — this favours the reference;
  - ``AutoTokenizer.from_pretrained`` -> a stock ``transformers.BertTokenizer`` over a synthetic vocabulary file,
    so tokenisation, ``DataCollatorWithPadding`` and the ``DataLoader`` are the stock code;
  - ``AutoModelForSequenceClassification.from_pretrained`` -> stock ``transformers.BertForSequenceClassification``
    from a ``BertConfig`` of the checkpoint's shape, random init (no hub); the subclass only adds a timestamp
    around ``save_pretrained`` (the round boundary);
  - ``transformers.AdamW`` (removed in transformers 5) -> ``torch.optim.AdamW`` with the removed class's defaults
    (betas (0.9, 0.999), eps 1e-6, weight_decay 0.0);
  - ``flwr`` -> a stub providing ``client.NumPyClient`` (base class only; the serverless script never starts a
    Flower server) and ``server.strategy.FedAvg`` (constructed, unused, :213-217);
  - ``evaluate.load("accuracy")`` -> a 10-line accuracy accumulator.

N > 1: the reference has no distributed path — rank 0 runs the script with ``NUM_CLIENTS = N``
(the same weak-scaled job the fedb200 arm runs on N GPUs), the other ranks exit.
"""
from __future__ import annotations

import ast
import contextlib
import hashlib
import io
import json
import os
import sys
import tempfile
import time
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.path.join(HERE, "_ref")
SCRIPT = os.path.join("src", "Serverlesscase", "serverless_IID_IMDB.py")


def _verify_unmodified() -> str:
    path = os.path.join(REF_ROOT, SCRIPT)
    sums = os.path.join(REF_ROOT, "SHA256SUMS")
    if not (os.path.exists(path) and os.path.exists(sums)) and os.path.isdir("/root/reference/src"):
        import subprocess                      # first use on a machine that has the reference mounted: stage it now
        subprocess.run([os.path.join(HERE, "install_reference.sh")], check=False, capture_output=True)
    if not (os.path.exists(path) and os.path.exists(sums)):
        raise FileNotFoundError("baseline/_ref is not staged (run baseline/install_reference.sh)")
    want = {l.split()[1]: l.split()[0] for l in open(sums) if l.strip()}
    got = hashlib.sha256(open(path, "rb").read()).hexdigest()
    if want.get(SCRIPT) != got:
        raise RuntimeError("reference script differs from its recorded checksum")
    live = os.path.join("/root/reference", SCRIPT)
    if os.path.exists(live) and hashlib.sha256(open(live, "rb").read()).hexdigest() != got:
        raise RuntimeError("staged reference script differs from /root/reference")
    return got


class _Rebind(ast.NodeTransformer):
    """Re-bind module-level constants (the reference's configuration surface)."""

    def __init__(self, values):
        self.values, self.hit = values, set()

    def visit_Module(self, node):
        for stmt in node.body:
            if isinstance(stmt, ast.Assign) and len(stmt.targets) == 1 and isinstance(stmt.targets[0], ast.Name) \
                    and stmt.targets[0].id in self.values:
                name = stmt.targets[0].id
                stmt.value = ast.parse(self.values[name], mode="eval").body
                self.hit.add(name)
        return node


# ---- offline stand-ins --------------------------------------------------------------------------------------------
class _Split:
    """One split of the stand-in DatasetDict: rows are materialised (and tokenised) only when selected."""

    def __init__(self, tokens, indices, fn=None, cols=None):
        self.tok, self.idx, self.fn, self.cols = tokens, indices, fn, cols

    def __len__(self):
        return len(self.idx)

    def shuffle(self, seed=None):
        import numpy as np
        return _Split(self.tok, self.idx[np.random.default_rng(seed).permutation(len(self.idx))], self.fn)

    def map(self, fn, batched=False):
        return _Split(self.tok, self.idx, fn)

    def select(self, indices):
        import numpy as np
        sub = self.idx[np.asarray(list(indices), dtype=np.int64)]
        texts, labels = [], []
        for i in sub:
            a, b = int(self.tok.offsets[i]) + 1, int(self.tok.offsets[i + 1]) - 1      # body without [CLS]/[SEP]
            texts.append(" ".join(f"w{t}" for t in self.tok.tokens[a:b]))
            labels.append(int(self.tok.labels[i]))
        cols = {"text": texts, "label": labels}
        if self.fn is not None:
            enc = self.fn({"text": texts})
            cols.update({k: list(v) for k, v in enc.items()})
        return _Split(self.tok, sub, None, cols)

    def remove_columns(self, name):
        cols = None if self.cols is None else {k: v for k, v in self.cols.items() if k != name}
        return _Split(self.tok, self.idx, self.fn, cols)

    def rename_column(self, old, new):
        cols = None if self.cols is None else {(new if k == old else k): v for k, v in self.cols.items()}
        return _Split(self.tok, self.idx, self.fn, cols)

    def __getitem__(self, i):
        return {k: v[i] for k, v in self.cols.items()}


class _DatasetDict(dict):
    def shuffle(self, seed=None):
        return _DatasetDict({k: v.shuffle(seed) for k, v in self.items()})

    def map(self, fn, batched=False):
        return _DatasetDict({k: v.map(fn, batched) for k, v in self.items()})

    def remove_columns(self, name):
        return _DatasetDict({k: v.remove_columns(name) for k, v in self.items()})

    def rename_column(self, old, new):
        return _DatasetDict({k: v.rename_column(old, new) for k, v in self.items()})


def _install_standins(preset: dict, seq_len: int, marks: list, ckpt_seconds: list, seed: int = 0):
    import numpy as np
    import torch
    import transformers
    from transformers import BertConfig, BertForSequenceClassification, BertTokenizer
    sys.path.insert(0, os.path.dirname(HERE))
    from fedb200.data.synthetic import make_synthetic

    vocab_size = preset["vocab_size"]
    corpus = {}

    def load_dataset(name, *a, **k):
        if name != "imdb":
            raise ValueError(f"offline stand-in only serves 'imdb', not {name!r}")
        if not corpus:
            for split in ("train", "test"):
                corpus[split] = make_synthetic("imdb", split, vocab_size, seq_len, None, 42, shuffled=False)
        d = _DatasetDict({s: _Split(t, np.arange(len(t))) for s, t in corpus.items()})
        d["unsupervised"] = _Split(corpus["train"], np.arange(0))
        return d

    ds_mod = types.ModuleType("datasets")
    ds_mod.load_dataset = load_dataset
    sys.modules["datasets"] = ds_mod

    ev_mod = types.ModuleType("evaluate")

    class _Accuracy:
        def __init__(self):
            self.c = self.n = 0

        def add_batch(self, predictions, references):
            self.c += int((predictions == references).sum())
            self.n += int(references.numel())

        def compute(self):
            return {"accuracy": self.c / max(self.n, 1)}
    ev_mod.load = lambda name: _Accuracy()
    sys.modules["evaluate"] = ev_mod

    fl = types.ModuleType("flwr")
    fl.client = types.ModuleType("flwr.client")
    fl.client.NumPyClient = type("NumPyClient", (), {})
    fl.server = types.ModuleType("flwr.server")
    fl.server.strategy = types.ModuleType("flwr.server.strategy")
    fl.server.strategy.FedAvg = type("FedAvg", (), {"__init__": lambda self, **kw: None})
    fl.server.ServerConfig = type("ServerConfig", (), {"__init__": lambda self, **kw: None})
    for n, m in (("flwr", fl), ("flwr.client", fl.client), ("flwr.server", fl.server), ("flwr.server.strategy", fl.server.strategy)):
        sys.modules[n] = m

    class AdamW(torch.optim.AdamW):           # transformers.AdamW defaults (class removed in transformers 5)
        def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0, **kw):
            super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
    setattr(sys.modules["transformers"], "AdamW", AdamW)

    words = [f"w{i}" for i in range(vocab_size)]
    words[0], words[100], words[101], words[102], words[103] = "[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"
    vocab = {w: i for i, w in enumerate(words)}

    def tok_from_pretrained(name, *a, **k):
        return BertTokenizer(vocab=vocab, do_lower_case=False, model_max_length=seq_len)
    transformers.AutoTokenizer.from_pretrained = staticmethod(tok_from_pretrained)

    class _TimedBert(BertForSequenceClassification):
        def save_pretrained(self, *a, **k):
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = super().save_pretrained(*a, **k)
            t1 = time.perf_counter()
            ckpt_seconds.append(t1 - t0)
            marks.append(t1)
            return out

    def model_from_pretrained(name, num_labels=2, **k):
        torch.manual_seed(seed)
        return _TimedBert(BertConfig(vocab_size=vocab_size, num_labels=num_labels, hidden_size=preset["hidden_size"],
                                     num_hidden_layers=preset["num_layers"], num_attention_heads=preset["num_heads"],
                                     intermediate_size=preset["intermediate_size"],
                                     max_position_embeddings=preset.get("max_position", 512)))
    transformers.AutoModelForSequenceClassification.from_pretrained = staticmethod(model_from_pretrained)


def run(args) -> dict:
    """Execute the reference script for W + K rounds; returns the bench JSON line (rank 0 only calls this)."""
    sha = _verify_unmodified()
    import torch
    device = os.environ.get("FEDB200_REF_DEVICE", "cuda")       # "cpu" only for the plumbing test of this file
    if device == "cuda" and not torch.cuda.is_available():
        return {"impl": "reference", "unavailable": "no CUDA device for the reference arm"}
    n = int(os.environ.get("WORLD_SIZE", "1"))
    K, W = args.steps, args.warmup
    from fedb200.config import MODEL_PRESETS
    preset = MODEL_PRESETS[args.model]
    seq_len = min(args.seq_len, preset.get("max_position", 512))
    marks, ckpt = [], []
    _install_standins(preset, seq_len, marks, ckpt)
    src = open(os.path.join(REF_ROOT, SCRIPT)).read()
    tree = ast.parse(src, filename=SCRIPT)
    values = {"DEVICE": ('torch.device("cuda", %d)' % int(os.environ.get("LOCAL_RANK", "0"))) if device == "cuda" else 'torch.device("cpu")',
              "CHECKPOINT": repr("dmis-lab/biobert-v1.1"), "NUM_CLIENTS": str(n * args.peers_per_gpu),
              "NUM_ROUNDS": str(W + K)}
    rb = _Rebind(values)
    tree = ast.fix_missing_locations(rb.visit(tree))
    assert rb.hit == set(values), f"constants not found: {set(values) - rb.hit}"
    code = compile(tree, SCRIPT, "exec")
    import bench as _bench
    sampler = _bench.ClockSampler(int(os.environ.get("LOCAL_RANK", "0")))
    cwd = os.getcwd()
    work = tempfile.mkdtemp(prefix="ref_arm_")
    os.chdir(work)
    sink = io.StringIO()
    glb = {"__name__": "__reference__", "__file__": SCRIPT}
    if device == "cuda":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    try:
        sampler.start()
        with contextlib.redirect_stdout(sink):
            exec(code, glb)
    finally:
        os.chdir(cwd)
    clocks = sampler.stop()
    assert len(marks) == W + K, (len(marks), W + K)
    total_s = marks[-1] - marks[W - 1]
    ck_s = sum(ckpt[W:])
    peers = n * args.peers_per_gpu
    samples = peers * args.train_per_client * K          # load_data(): 100 random train rows per client visit (:58)
    accs = glb.get("global_model_Accuracies", [])
    excl = total_s - ck_s
    value = samples / excl
    h2d = peers * (args.train_per_client + args.test_per_client) * seq_len * 3 * 8      # upper bound: padded int64 x3
    return {
        "metric": "train_samples_per_s (FL round = local epoch + local eval + aggregate + global eval)",
        "value": round(value, 2), "unit": "samples/s", "n_gpus": n, "steps": K, "warmup": W,
        "ms_per_step": round(excl / K * 1e3, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": round(value / _bench.PUBLISHED_TRAIN_SAMPLES_PER_S, 2), "dtype": "fp32", "data": "synthetic",
        "impl": "reference",
        "config": {"script": SCRIPT, "sha256": sha, "model": "transformers BertForSequenceClassification, BioBERT shape, random init",
                   "rebound_constants": values, "peers": peers, "batch_per_peer": 32, "seq_len": seq_len,
                   "train_per_peer_per_round": args.train_per_client, "test_per_peer_per_round": args.test_per_client,
                   "parallelism": "single process, clients visited sequentially on cuda:0 (the reference has no multi-GPU path)",
                   "standins": ["datasets.load_dataset -> synthetic IMDB-shape corpus (lazy map)", "AutoTokenizer -> BertTokenizer(synthetic vocab)",
                                "from_pretrained -> BertConfig random init", "transformers.AdamW -> torch.optim.AdamW(eps 1e-6, wd 0)",
                                "flwr stub (NumPyClient base class)", "evaluate.load('accuracy') stub"],
                   "checkpoint": "the script's per-round save_pretrained is timed separately and EXCLUDED from value "
                                 "(the fedb200 arm does not checkpoint in its timed region)",
                   "ms_per_step_incl_checkpoint": round(total_s / K * 1e3, 3),
                   "global_accuracy_per_round": [round(float(a), 4) for a in accs]},
        "clocks": clocks,
        "e2e": {"value": round(value, 2), "unit": "samples/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(peers * 433_364_128 + 8 * peers),
                "api": "module execution of the unmodified script (host batches -> .to(DEVICE) every step, full-model D2H per client)"},
        "gpu_launches": 0,
    }


def main(args) -> None:
    if int(os.environ.get("RANK", "0")) != 0:
        return
    try:
        line = run(args)
    except Exception as e:      # the contract: print an 'unavailable' line and exit 0
        line = {"impl": "reference", "unavailable": f"{type(e).__name__}: {str(e)[:300]}"}
    print(json.dumps(line))
