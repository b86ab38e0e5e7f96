"""Relative error of the attention forward against an fp32 PyTorch reference, per sequence (no dropout).

    FEDB200_ATTN_FWD=2|4 python bench/attn_fwd_check.py
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from fedb200.ops import cuda_ops
heads, H = 4, 256
lens = [512, 448, 400, 385, 384, 300, 257, 256, 200, 160, 129, 128, 100, 65, 64, 33, 32, 7, 1]
T, B = sum(lens), len(lens)
torch.manual_seed(1)
cu = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32, device="cuda")
out = {"variant": os.environ.get("FEDB200_ATTN_FWD", "default")}
for amp in (0.5, 3.0):
    qkv = (torch.randn(T, 3 * H, device="cuda") * amp).bfloat16()
    ctx = torch.zeros(T, H, dtype=torch.bfloat16, device="cuda"); lse = torch.zeros(heads * T, device="cuda")
    cuda_ops.attn_fwd(qkv, ctx, lse, cu, B, heads, max(lens), 0.125, 7, 0.0)
    torch.cuda.synchronize()
    errs, lerr = [], []
    f = qkv.float()
    for b in range(B):
        s, e = int(cu[b]), int(cu[b + 1])
        num = den = 0.0
        for h in range(heads):
            q, k, v = f[s:e, h * 64:(h + 1) * 64], f[s:e, H + h * 64:H + (h + 1) * 64], f[s:e, 2 * H + h * 64:2 * H + (h + 1) * 64]
            sc = (q @ k.T) * 0.125
            ref = torch.softmax(sc, -1) @ v
            got = ctx[s:e, h * 64:(h + 1) * 64].float()
            num += float((got - ref).pow(2).sum()); den += float(ref.pow(2).sum())
            l2 = torch.logsumexp(sc, -1) * 1.4426950408889634
            lerr.append(float((lse[h * T + s:h * T + e] - l2).abs().max()))
        errs.append(round((num / max(den, 1e-30)) ** 0.5, 5))
    out[f"rel_err_by_seq_amp{amp}"] = errs
    out[f"lse_max_err_amp{amp}"] = round(max(lerr), 5)
print(json.dumps(out))
