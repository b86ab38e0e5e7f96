"""What bounds the attention forward?  Runs the traced four-thread kernel with parts switched off (results are then wrong —
timing only) and with one CTA per SM, on the IMDB-shape batch of bench/attn_fwd_time.py.

    FEDB200_ATTN_FWD=4 python bench/attn_diag.py
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from fedb200.data import make_synthetic
from fedb200.ops import cuda_ops, _lib
heads, H = 12, 768
ds = make_synthetic("imdb", "train", 28996, 512, rows=64)
lens = sorted((int(x) for x in ds.lengths()[:32]), reverse=True)
T, B = sum(lens), len(lens)
cu = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32, device="cuda")
torch.manual_seed(0)
qkv = (torch.randn(T, 3 * H, device="cuda") * 0.5).bfloat16()
ctx = torch.zeros(T, H, dtype=torch.bfloat16, device="cuda"); lse = torch.zeros(heads * T, device="cuda")
bits = torch.zeros(heads * T * 16, dtype=torch.int32, device="cuda")
nq = (max(lens) + 127) // 128
trace = torch.zeros(nq * heads * B * 32, dtype=torch.int64, device="cuda")
_lib.call("fedb200_attn_set_trace", trace.data_ptr())
names = {0: "full kernel", 1: "no exp2", 2: "no TMEM loads in the sweep", 4: "no P stores", 7: "none of the three"}
for pad in (0, 48 * 1024):
    for p in (0.0, 0.1):
        for diag in (0, 1, 2, 4, 7):
            _lib.call("fedb200_attn_set_diag", diag | 8, pad)
            ts = []
            for it in range(8):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); cuda_ops.attn_fwd(qkv, ctx, lse, cu, B, heads, max(lens), 0.125, 7, p, drop_bits=bits if p else None); e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1) * 1e3)
            t = trace.cpu().numpy().reshape(-1, 32)
            t = t[t[:, 30] > 0]
            sweep, swait, owait = [], [], []
            for row in t:
                for j in range(int(row[29])):
                    b = 3 + 5 * j
                    sweep.append(row[b + 1] - row[b + 3]); owait.append(row[b + 3] - row[b])
                    if j > 0:
                        swait.append(row[b] - row[b - 5 + 2])
            t2 = t[t[:, 29] >= 2]            # CTAs with a second tile: its sweep is stamped in detail
            det = [int(np.mean(t2[:, a] - t2[:, b])) for a, b in ((23, 3 + 5 + 3), (24, 23), (25, 24), (26, 25), (27, 26), (3 + 5 + 1, 27))]
            print(f"CTAs/SM {1 if pad else 2} dropout {p} {names[diag]:28s}: {np.median(ts):6.1f} us  sweep+xchg {np.mean(sweep):5.0f}  "
                  f"wait S {np.mean(swait):4.0f}  wait O {np.mean(owait):4.0f} clk | tile 1: ld0 {det[0]} math0 {det[1]} ld1 {det[2]} math1 {det[3]} "
                  f"sts {det[4]} bar {det[5]}")
_lib.call("fedb200_attn_set_diag", 8, 0)
_lib.call("fedb200_attn_set_trace", None)
