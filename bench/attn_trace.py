"""Phase timeline of the attention forward: clock64() stamps per CTA (fedb200_attn_set_trace) on an IMDB-shape batch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from fedb200.data import make_synthetic
from fedb200.ops import cuda_ops, _lib
heads, H = 12, 768
ds = make_synthetic("imdb", "train", 28996, 512, rows=64)
lens = [int(x) for x in ds.lengths()[:32]]
T, B = sum(lens), len(lens)
cu = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32, device="cuda")
qkv = (torch.randn(T, 3 * H, device="cuda") * 0.5).bfloat16()
ctx = torch.zeros(T, H, dtype=torch.bfloat16, device="cuda"); lse = torch.zeros(heads * T, device="cuda")
bits = torch.zeros(heads * T * 16, dtype=torch.int32, device="cuda")
nq = (max(lens) + 127) // 128
trace = torch.zeros(nq * heads * B * 32, dtype=torch.int64, device="cuda")
for p in (0.1, 0.0):
    for it in range(3):
        cuda_ops.attn_fwd(qkv, ctx, lse, cu, B, heads, max(lens), 0.125, 7, p, drop_bits=bits if p else None)
    torch.cuda.synchronize()
    trace.zero_()
    _lib.call("fedb200_attn_set_trace", trace.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); cuda_ops.attn_fwd(qkv, ctx, lse, cu, B, heads, max(lens), 0.125, 7, p, drop_bits=bits if p else None); e1.record()
    torch.cuda.synchronize()
    _lib.call("fedb200_attn_set_trace", None)
    t = trace.cpu().numpy().reshape(-1, 32)
    t = t[t[:, 30] > 0]
    t0 = t[:, 0].min()
    dur = t[:, 30] - t[:, 0]
    nk = t[:, 29]
    print(f"dropout {p}: kernel {e0.elapsed_time(e1)*1e3:.1f} us, {len(t)} CTAs, span {(t[:,30].max()-t0)} clk, CTA duration mean {dur.mean():.0f} clk "
          f"(n_kv 1..4: {[int(dur[nk==k].mean()) if (nk==k).any() else 0 for k in (1,2,3,4)]})")
    setup = t[:, 2] - t[:, 0]
    first = t[:, 3] - t[:, 2]
    print(f"  launch->after pdl/setup {setup.mean():.0f} clk, setup->first s_full {first.mean():.0f} clk")
    ph = {"arrive->s_full(next)": [], "pass1": [], "wait_o(after pass1)": [], "fold": [], "pass2": []}
    if os.environ.get("FEDB200_ATTN_FWD", "4") == "4":
        # four-thread kernel: stamps 3+5j after s_full, 4+5j after max pass + exchange, 6+5j after rescale decision / o_full wait,
        # 5+5j after the probability pass and the p_ready arrive
        ph = {"arrive->s_full(next)": [], "pass1+xchg": [], "rescale/wait_o": [], "pass2": []}
        for row in t:
            for j in range(int(row[29])):
                b = 3 + 5 * j
                ph["pass1+xchg"].append(row[b + 1] - row[b])
                ph["rescale/wait_o"].append(row[b + 3] - row[b + 1])
                ph["pass2"].append(row[b + 2] - row[b + 3])
                if j > 0:
                    ph["arrive->s_full(next)"].append(row[b] - row[b - 5 + 2])
        print("  per kv tile (clk):", {k: int(np.mean(v)) for k, v in ph.items() if v}, "tiles", len(ph["pass1+xchg"]))
        t = t[:0]
    for row in t:
        n = int(row[29])
        for j in range(n):
            b = 3 + 5 * j
            ph["pass1"].append(row[b + 1] - row[b])
            if j > 0:
                pb = 3 + 5 * (j - 1)
                ph["arrive->s_full(next)"].append(row[b] - row[pb + 2])
                ph["wait_o(after pass1)"].append(row[pb + 3] - row[b + 1])
                ph["fold"].append(row[pb + 4] - row[pb + 3])
                ph["pass2"].append(row[b + 2] - row[pb + 4])
            else:
                ph["pass2"].append(row[b + 2] - row[b + 1])
    if len(t):
        print("  per kv tile (clk):", {k: int(np.mean(v)) for k, v in ph.items() if v}, "tiles", len(ph["pass1"]))
    t = trace.cpu().numpy().reshape(-1, 32)
    t = t[t[:, 30] > 0]
    # CTAs per SM and the busy span per SM
    sm = t[:, 31]
    per_sm = [(t[sm == s, 30].max() - t[sm == s, 0].min(), (sm == s).sum(), dur[sm == s].sum()) for s in np.unique(sm)]
    print("  per SM: span mean %.0f clk, CTAs %.1f, sum of CTA durations %.0f clk" % tuple(np.mean(per_sm, axis=0)))
