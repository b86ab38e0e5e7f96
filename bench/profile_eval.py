"""Eval (forward-only) step for ncu launch lists."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from fedb200.config import FedConfig, ModelConfig
from fedb200.data import make_synthetic
from fedb200.data.dataset import collate_packed
from fedb200.models.fused_bert import FusedRunner
fed = FedConfig(); fed.model = ModelConfig.preset("biobert", num_labels=2)
run = FusedRunner(fed.model, fed, "cuda", seed=0)
ds = make_synthetic("imdb", "train", fed.model.vocab_size, 512, rows=64)
b = collate_packed(ds, np.arange(32)); b = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in b.items()}
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    run.eval_step(b)
torch.cuda.synchronize()
print("done")
