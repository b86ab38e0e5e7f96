"""Cost of ``--deterministic`` (ordered reductions, csrc/determinism.cu) on the flagship training step.

    python bench/det_overhead.py          # BioBERT shape, 32 sequences of the synthetic IMDB-shape corpus, CUDA-event times

Prints one JSON line: ms per training step in the default (fp32 atomics) and the reproducible mode, and whether two
reproducible runs ended with identical bytes.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from fedb200.config import FedConfig, ModelConfig
from fedb200.data import make_synthetic
from fedb200.data.dataset import collate_packed
from fedb200.models.fused_bert import FusedRunner

model = os.environ.get("FEDB200_DET_MODEL", "biobert")
fed = FedConfig()
fed.model = ModelConfig.preset(model, num_labels=2)
ds = make_synthetic("imdb", "train", fed.model.vocab_size, 512, rows=128)
batches = [collate_packed(ds, np.arange(i * 32, (i + 1) * 32), longest_first=True) for i in range(4)]
batches = [{k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in b.items()} for b in batches]


def run(det: bool, steps: int = 12):
    fed.deterministic = det
    r = FusedRunner(fed.model, fed, "cuda", seed=0)
    for b in batches:                       # warm-up
        r.train_step(b)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    torch.cuda.synchronize()
    for i in range(steps):
        flush.zero_()                       # L2 flush between timed steps
        ev[i][0].record()
        r.train_step(batches[i % 4])
        ev[i][1].record()
    torch.cuda.synchronize()
    ms = float(np.median([a.elapsed_time(b) for a, b in ev]))
    w = r.flat.clone()
    if det:
        r.set_deterministic(False)
    del r
    return ms, w


ms0, _ = run(False)
ms1, w1 = run(True)
ms2, w2 = run(True)
print(json.dumps({"model": model, "tokens_per_step": int(batches[0]["num_tokens"]), "ms_step_default": round(ms0, 3),
                  "ms_step_deterministic": round(ms1, 3), "overhead": round(ms1 / ms0 - 1, 4),
                  "two_deterministic_runs_identical": bool(torch.equal(w1, w2)),
                  "differing_elements": int((w1 != w2).sum().item())}))
