"""SM clock / power distribution while training steps run back to back (is the step power-capped?)."""
import sys, os, subprocess, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from fedb200.config import FedConfig, ModelConfig
from fedb200.data import make_synthetic
from fedb200.data.dataset import collate_packed
from fedb200.models.fused_bert import FusedRunner

fed = FedConfig(); fed.model = ModelConfig.preset("biobert", num_labels=2)
run = FusedRunner(fed.model, fed, "cuda", seed=0)
ds = make_synthetic("imdb", "train", fed.model.vocab_size, 512, rows=64)
b32 = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in collate_packed(ds, np.arange(32)).items()}
for _ in range(3):
    run.train_step(b32)
torch.cuda.synchronize()
p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.active", "--format=csv,noheader,nounits", "-lms", "20", "-i", "0"],
                     stdout=subprocess.PIPE, text=True)
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.time(); n = 0
ev0.record()
while time.time() - t0 < 3.0:
    for _ in range(10):
        run.train_step(b32)
    n += 10
    torch.cuda.synchronize()
ev1.record(); torch.cuda.synchronize()
p.terminate()
rows = [l.split(",") for l in p.stdout.read().strip().split("\n") if l.count(",") == 2]
rows = rows[len(rows) // 4:]
clk = np.array([float(r[0]) for r in rows]); pw = np.array([float(r[1]) for r in rows])
capped = np.mean([int(r[2].strip(), 16) & 4 != 0 for r in rows])
print(f"train step {ev0.elapsed_time(ev1) / n:.3f} ms over {n} steps; SM clock MHz min/median/max {clk.min():.0f}/{np.median(clk):.0f}/{clk.max():.0f}; "
      f"power W median {np.median(pw):.0f} max {pw.max():.0f}; sw_power_cap active in {100 * capped:.0f}% of {len(rows)} samples")
