"""NVLS (multimem) FedAvg reduce vs NCCL and vs the P2P two-shot kernel: correctness + device time.  torchrun, one process per GPU."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from fedb200.parallel import init_distributed, make_comm

rank, world, local = init_distributed(torch.device("cuda"))
dev = torch.device("cuda", local)
comm = make_comm("peer", dev, rank, world)
ok = comm.nvls_available()
if rank == 0:
    print("nvls available:", ok)
if ok:
    torch.manual_seed(7 + rank)
    n = int(os.environ.get("NVLS_N", 108_312_064))
    acc = comm.alloc("nv_acc", n, multicast=True)
    out = comm.alloc("nv_out", n, multicast=True)
    x = torch.randn(n, device=dev)
    for it, n_use in enumerate((1_000_003, n)):
        acc[:n_use].copy_(x[:n_use] * (rank + 1))
        torch.cuda.synchronize(); dist.barrier()
        comm.weighted_mean(acc[:n_use], float(rank + 1), out[:n_use], nvls=True)
        ref = (x[:n_use] * (rank + 1)).clone()
        dist.all_reduce(ref)
        ref /= sum(range(1, world + 1))
        torch.cuda.synchronize()
        comm.check_errors()
        err = (out[:n_use] - ref).abs().max().item()
        assert err < 1e-5, f"nvls mismatch {err}"
        if rank == 0:
            print(f"nvls reduce n={n_use}: max |diff| vs NCCL {err:.2e}")
    def timed(fn, iters=10):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(iters):
            torch.cuda.synchronize(); dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        t = torch.tensor([float(np.median(ts))], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()
    t_nvls = timed(lambda: comm.weighted_mean(acc, 1.0, out, nvls=True))
    t_nvls32 = timed(lambda: comm.weighted_mean(acc, 1.0, out, nvls=True, max_ctas=32))
    t_p2p = timed(lambda: comm.weighted_mean(acc, 1.0, out))
    buf = acc.clone()
    t_nccl = timed(lambda: (dist.all_reduce(buf), buf.mul_(1.0 / world)))
    if rank == 0:
        print(f"NVLS_RESULT world={world} n={n}: nvls {t_nvls:.3f} ms (32 CTAs: {t_nvls32:.3f}), p2p two-shot {t_p2p:.3f} ms, NCCL all_reduce+scale {t_nccl:.3f} ms")
comm.close()
dist.destroy_process_group()
