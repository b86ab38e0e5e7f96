"""Component timing of one FL round at N=1 (device time, CUDA events, L2 flushed between iterations)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from fedb200.config import FedConfig, ModelConfig
from fedb200.data import make_synthetic
from fedb200.data.dataset import collate_packed
from fedb200.models.fused_bert import FusedRunner
from fedb200.ops import cuda_ops as ops
from fedb200.utils.timers import cuda_time_ms

fed = FedConfig()
fed.model = ModelConfig.preset("biobert", num_labels=2)
run = FusedRunner(fed.model, fed, "cuda", seed=0)
ds = make_synthetic("imdb", "train", fed.model.vocab_size, 512, rows=64)
dev = lambda b: {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in b.items()}
b32, b4 = dev(collate_packed(ds, np.arange(32))), dev(collate_packed(ds, np.arange(32, 36)))
D = run.flat.numel()
other = torch.randn(D, device="cuda")
out = torch.empty(D, device="cuda")
nch = (D * 4 + 4095) // 4096
leaves = torch.empty(nch, 32, dtype=torch.uint8, device="cuda")
scratch = torch.empty(nch // 2 + 2, 32, dtype=torch.uint8, device="cuda")
root = torch.empty(32, dtype=torch.uint8, device="cuda")
t = lambda f, **kw: cuda_time_ms(f, warmup=2, iters=5, **kw)
print(f"tokens b32={b32['num_tokens']} b4={b4['num_tokens']}  D={D}")
print(f"train_step b32      {t(lambda: run.train_step(b32)):8.3f} ms")
print(f"train_step b4       {t(lambda: run.train_step(b4)):8.3f} ms")
print(f"eval_step  b32      {t(lambda: run.eval_step(b32)):8.3f} ms")
print(f"eval_step  b4       {t(lambda: run.eval_step(b4)):8.3f} ms")
print(f"adamw (D)           {t(lambda: run.opt.step()):8.3f} ms   roofline {D*34/6555e6:.3f} ms")
print(f"sync_shadow cast    {t(run.sync_shadow):8.3f} ms   roofline {D*6/6555e6:.3f} ms")
print(f"load_flat copy      {t(lambda: run.load_flat(other)):8.3f} ms   roofline {D*8/6555e6:.3f} ms")
print(f"opt.reset (2 memset){t(run.opt.reset):8.3f} ms")
print(f"sha256 leaves+copy  {t(lambda: ops.sha256_leaves(run.flat, out, 4096, leaves)):8.3f} ms   ({D*4/1e6:.0f} MB)")
print(f"sha256 leaves only  {t(lambda: ops.sha256_leaves(run.flat, None, 4096, leaves)):8.3f} ms")
print(f"merkle root         {t(lambda: ops.merkle_root(leaves, scratch, root)):8.3f} ms")
print(f"mix_n 1 src         {t(lambda: ops.mix_n([run.flat], [1.0], out)):8.3f} ms   roofline {D*8/6555e6:.3f} ms")
print(f"mix_n 8 src         {t(lambda: ops.mix_n([run.flat, other] * 4, [0.125] * 8, out)):8.3f} ms")
d16 = torch.randn(8, D // 8 * 8, device="cuda").bfloat16()
print(f"gram 8 x D bf16     {t(lambda: ops.gram(d16)):8.3f} ms   roofline {8*D*2/6555e6:.3f} ms")
d64 = torch.randn(64, D // 64 * 8, device="cuda").bfloat16()
print(f"gram 64 x D/8 bf16  {t(lambda: ops.gram(d64)):8.3f} ms   roofline {64*(D//64*8)*2/6555e6:.3f} ms")
