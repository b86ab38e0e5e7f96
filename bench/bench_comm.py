"""Fused collective kernels vs their NVLink roofline and vs NCCL (run under torchrun, one process per GPU).

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 bench/bench_comm.py [--model biobert]

Every number is device-timed (CUDA events) and reduced as the MAX over ranks.  Roofline = bytes that must cross
NVLink per GPU per direction / 770 GB/s (measured peer-copy bandwidth of this pool, B200_PROFILING.md)."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from fedb200.config import ModelConfig
from fedb200.models.arena import param_layout
from fedb200.models.factory import build_model
from fedb200.parallel import init_distributed, make_comm
from fedb200.ops import cuda_ops

LINK = 770e9


def timed(fn, dev, iters=10, warmup=3):
    for _ in range(warmup):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    t = torch.tensor([float(np.median(ts))], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="biobert")
    a = ap.parse_args()
    rank, world, local = init_distributed(torch.device("cuda"))
    dev = torch.device("cuda", local)
    comm = make_comm("peer", dev, rank, world)
    with torch.device("meta"):
        m = build_model(ModelConfig.preset(a.model, num_labels=2), "meta")
    D = sum(v.numel() for v in m.state_dict().values())
    D = (D + 255) // 256 * 256
    G = world
    rows = []
    acc = comm.alloc("b_acc", D); out = comm.alloc("b_out", D); shadow = comm.alloc("b_shadow", D, torch.bfloat16)
    acc.copy_(torch.randn(D, device=dev))
    # ---- K1 fused FedAvg reduce (two-shot, + scale + bf16 shadow + broadcast) vs NCCL all_reduce + ATen --------------
    ms = timed(lambda: comm.weighted_mean(acc, 1.0, out, shadow), dev)
    bytes_dir = 2 * (G - 1) / G * D * 4 + (G - 1) / G * D * 2        # reduce-scatter reads + all-gather pushes (fp32 + bf16 shadow)
    rows.append(("fedavg_reduce (fp32 in, fp32 + bf16 out)", ms, bytes_dir / LINK * 1e3))
    ms = timed(lambda: comm.weighted_mean(acc, 1.0, out), dev)
    bytes_dir = 2 * (G - 1) / G * D * 4
    rows.append(("fedavg_reduce (fp32 in/out)", ms, bytes_dir / LINK * 1e3))
    buf = acc.clone()
    def nccl():
        dist.all_reduce(buf); buf.mul_(1.0 / G); shadow.copy_(buf)
    rows.append(("NCCL all_reduce + mul + bf16 cast (baseline)", timed(nccl, dev), bytes_dir / LINK * 1e3))
    # ---- K4 all-gather-fused Gram (tcgen05 + TMA over peer memory) vs all_gather + matmul -----------------------------
    deltas = comm.delta_buffer(1, D, torch.bfloat16)
    deltas[:, :D].copy_((torch.randn(1, D, device=dev) * 0.01).bfloat16())
    ms = timed(lambda: comm.gram(deltas, G), dev)
    remote = G * D * (G - 1) / (G * G) * 2                               # remote bytes read per GPU
    local_hbm = G * D / G * 2 / 6555e9 * 1e3
    rows.append((f"cosine_gram {G} peers (bf16, all-gather fused)", ms, max(remote / LINK * 1e3, local_hbm)))
    gl = [torch.empty_like(deltas) for _ in range(G)]
    def nccl_gram():
        dist.all_gather(gl, deltas)
        x = torch.cat(gl, 0)
        return x.float() @ x.float().T
    rows.append(("NCCL all_gather + fp32 matmul (baseline)", timed(nccl_gram, dev, iters=5), (G - 1) * D * 2 / LINK * 1e3))
    # ---- K2 routed gossip mix on a ring (2 neighbours + self) ------------------------------------------------------------------
    slots = comm.alloc("b_slots", D).view(1, D); mixed = comm.alloc("b_mixed", D).view(1, D)
    slots.copy_(torch.randn(1, D, device=dev))
    W = np.zeros((G, G)); 
    for i in range(G):
        W[i, i] = W[i, (i + 1) % G] = W[i, (i - 1) % G] = 1.0 / 3.0
    if G == 2:
        W = np.full((2, 2), 0.5)
    deg = 2 if G > 2 else 1
    ms = timed(lambda: comm.mix_rows(slots, W[rank:rank + 1], mixed, G), dev)
    rows.append((f"gossip_mix ring (deg {deg}, incl. 2 device barriers)", ms, deg * D * 4 / LINK * 1e3))
    # ---- K6 SHA-256 fused into the publish copy (local hash-bound; peer destination) --------------------------------------------
    leaves = torch.empty((D * 4 + 4095) // 4096, 32, dtype=torch.uint8, device=dev)
    nxt = (rank + 1) % G
    dst_ptr = comm.peer_ptr(out, nxt)
    ms = timed(lambda: cuda_ops.sha256_leaves(acc, None, 4096, leaves, dst_ptr=dst_ptr), dev)
    rows.append(("sha256_send to peer (433 MB hashed while copied over NVLink)", ms, D * 4 / LINK * 1e3))
    comm.check_errors()
    if rank == 0:
        res = {"n_gpus": G, "model": a.model, "params": D, "link_GBps_per_dir": LINK / 1e9,
               "rows": [{"kernel": n, "ms": round(t, 4), "roofline_ms": round(r, 4), "fraction_of_roofline": round(r / t, 3)} for n, t, r in rows]}
        print(json.dumps(res))
        print(f"\n| kernel ({G} x B200, D = {D/1e6:.1f} M) | time ms | roofline ms | fraction |\n|---|---|---|---|")
        for n, t, r in rows:
            print(f"| {n} | {t:.3f} | {r:.3f} | {r / t:.2f} |")
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
