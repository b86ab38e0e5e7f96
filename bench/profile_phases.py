"""Device time per FedRuntime phase of the bench configuration (N = 1), through the public round_serverless()."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench as B
from fedb200.runtime import FedRuntime

args = argparse.Namespace(steps=5, warmup=3, model="biobert", seq_len=512, batch_size=32, train_per_client=100,
                          test_per_client=100, peers_per_gpu=1, comm="auto", no_ledger=False)
cfg = B.make_cfg(args, 1)
rt = FedRuntime(cfg, echo=False)
for r in range(3):
    rt.round_serverless(r)
torch.cuda.synchronize()
rt.timer.reset()
K = 5
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ev0.record()
for r in range(3, 3 + K):
    rt.round_serverless(r)
ev1.record()
torch.cuda.synchronize()
tot = ev0.elapsed_time(ev1) / K
ms = rt.timer.device_ms()
print(f"round {tot:.2f} ms device")
for k, v in sorted(ms.items(), key=lambda x: -x[1]):
    print(f"  {k:18s} {v / K:8.3f} ms   wall {rt.timer.wall[k] / K:8.3f} ms")
print(f"  {'(unattributed)':18s} {tot - sum(ms.values()) / K:8.3f} ms")
