"""Device time of the attention backward (delta + main + finish kernels) on the IMDB-shape batch of attn_fwd_time.py.

    FEDB200_ATTN_BWD_PT=smem|tmem python bench/attn_bwd_time.py
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from fedb200.data import make_synthetic
from fedb200.ops import cuda_ops
heads, H = 12, 768
ds = make_synthetic("imdb", "train", 28996, 512, rows=64)
lens = sorted((int(x) for x in ds.lengths()[:32]), reverse=True)
T, B = sum(lens), len(lens)
torch.manual_seed(0)
cu = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32, device="cuda")
qkv = (torch.randn(T, 3 * H, device="cuda") * 0.5).bfloat16()
dctx = (torch.randn(T, H, device="cuda") * 0.5).bfloat16()
ctx = torch.zeros(T, H, dtype=torch.bfloat16, device="cuda"); lse = torch.zeros(heads * T, device="cuda")
bits = torch.zeros(heads * T * 16, dtype=torch.int32, device="cuda")
dqkv = torch.zeros(T, 3 * H, dtype=torch.bfloat16, device="cuda")
dq_acc = torch.empty((max(lens) + 127) // 128 * T, H, device="cuda")
delta = torch.zeros(heads * T, device="cuda")
dbias = torch.zeros(3 * H, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
out = {"variant": os.environ.get("FEDB200_ATTN_BWD_PT", "default"), "tokens": T}
for p in (0.1, 0.0):
    cuda_ops.attn_fwd(qkv, ctx, lse, cu, B, heads, max(lens), 0.125, 7, p, drop_bits=bits if p else None)
    ts = []
    for it in range(25):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        cuda_ops.attn_bwd(qkv, ctx, dctx, lse, delta, dqkv, dq_acc, cu, B, heads, max(lens), 0.125, 7, p, dbias=dbias,
                          drop_bits=bits if p else None)
        e1.record()
        torch.cuda.synchronize()
        if it >= 5:
            ts.append(e0.elapsed_time(e1) * 1e3)
    out[f"us_dropout_{p}"] = round(float(np.median(ts)), 2)
    out[f"dqkv_checksum_{p}"] = round(float(dqkv.float().abs().sum()), 1)
print(json.dumps(out))
