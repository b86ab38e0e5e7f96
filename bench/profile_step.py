"""One fused BERT-base train step on an IMDB-shape batch: CPU launch time vs GPU time, and a marker for ncu."""
import argparse, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from fedb200.config import FedConfig, ModelConfig
from fedb200.data import make_synthetic
from fedb200.data.dataset import collate_packed
from fedb200.models.fused_bert import FusedRunner
from fedb200.ops import _lib

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--model", default="biobert")
ap.add_argument("--dropout", type=float, default=0.1)
ap.add_argument("--full-len", action="store_true", help="every sequence at 512 tokens")
ap.add_argument("--batch", type=int, default=32)
a = ap.parse_args()
fed = FedConfig()
fed.model = ModelConfig.preset(a.model, num_labels=2, hidden_dropout=a.dropout, attention_dropout=a.dropout)
run = FusedRunner(fed.model, fed, "cuda", seed=0)
ds = make_synthetic("imdb", "train", fed.model.vocab_size, 512, rows=64)
if a.full_len:
    from fedb200.data.dataset import TokenDataset
    rng = np.random.default_rng(0)
    ds = TokenDataset.from_sequences([rng.integers(1000, 20000, size=512) for _ in range(32)], rng.integers(0, 2, size=32), 2)
batch = collate_packed(ds, np.arange(a.batch), longest_first=os.environ.get('FEDB200_SORT', '1') == '1')
dev_batch = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
print("tokens", batch["num_tokens"], "max_len", batch["max_len"])
for _ in range(a.warmup):
    run.train_step(dev_batch)
torch.cuda.synchronize()
l0 = _lib.LAUNCHES
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for _ in range(a.steps):
    run.train_step(dev_batch)
e1.record()
t_cpu = time.perf_counter() - t0
torch.cuda.synchronize()
gpu_ms = e0.elapsed_time(e1) / a.steps
T = batch["num_tokens"]
c = fed.model
flops = 6 * T * (c.num_layers * (4 * c.hidden_size ** 2 + 2 * c.hidden_size * c.intermediate_size)) \
    + 3 * 4 * c.num_layers * c.hidden_size * sum(int(l) ** 2 for l in np.diff(batch["cu_seqlens"].numpy()))
print(f"train step: gpu {gpu_ms:.3f} ms, cpu-issue {t_cpu / a.steps * 1e3:.3f} ms, launches/step {(_lib.LAUNCHES - l0) / a.steps:.0f}, "
      f"{flops / gpu_ms / 1e9:.1f} TFLOP/s model flops, {T / gpu_ms * 1e3:.0f} tokens/s")
# eval step
for _ in range(2):
    run.eval_step(dev_batch)
torch.cuda.synchronize()
e0.record()
for _ in range(a.steps):
    run.eval_step(dev_batch)
e1.record()
torch.cuda.synchronize()
print(f"eval step: gpu {e0.elapsed_time(e1) / a.steps:.3f} ms")

import ctypes as _C
_h, _m = _C.c_ulonglong(0), _C.c_ulonglong(0)
_lib.lib().fedb200_tmap_cache_stats(_C.byref(_h), _C.byref(_m))
print(f"tensor-map cache: {_h.value} hits, {_m.value} encodes (FEDB200_TMAP_CACHE={os.environ.get('FEDB200_TMAP_CACHE', '1')})")
