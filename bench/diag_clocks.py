"""SM clock / power while one kernel class runs back to back for ~1.5 s (is the GEMM loop power-capped?)."""
import sys, os, subprocess, time, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedb200.ops import cuda_ops as ops

T = 9187
dev = "cuda"
x = torch.randn(T, 768, device=dev).bfloat16(); w1 = (torch.randn(3072, 768, device=dev) * 0.02).bfloat16()
b1 = torch.randn(3072, device=dev); z = torch.empty(T, 3072, device=dev, dtype=torch.bfloat16)
big = torch.empty(1 << 28, device=dev, dtype=torch.float32)


def sample(tag, fn, secs=1.5):
    p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.active", "--format=csv,noheader", "-lms", "50", "-i", "0"],
                         stdout=subprocess.PIPE, text=True)
    t0 = time.time(); n = 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); ev0.record()
    while time.time() - t0 < secs:
        for _ in range(50):
            fn()
        n += 50
        torch.cuda.synchronize()
    ev1.record(); torch.cuda.synchronize()
    p.terminate()
    lines = [l.strip() for l in p.stdout.read().strip().split("\n") if l.strip()]
    mid = lines[len(lines) // 3:]
    print(f"{tag}: {ev0.elapsed_time(ev1) * 1000 / n:.1f} us/iter  samples (last 2/3): " + " | ".join(mid[:: max(1, len(mid) // 6)]))


sample("gemm ffn-up discard", lambda: ops.gemm(0, x, w1, z, T, 3072, 768, epi=6, bias=b1))
sample("gemm ffn-up bias", lambda: ops.gemm(0, x, w1, z, T, 3072, 768, epi=ops.EPI_BIAS, bias=b1))
sample("torch.matmul", lambda: torch.matmul(x, w1.T, out=z))
sample("memset 1 GiB", lambda: big.zero_(), 1.0)
