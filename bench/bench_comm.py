"""Fused collective kernels vs their NVLink roofline and vs NCCL (run under torchrun, one process per GPU).

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 bench/bench_comm.py [--model biobert]

Every number is device-timed (CUDA events) and reduced as the MAX over ranks.  The link bandwidth used as the roofline
denominator is MEASURED here, first: every rank pulls D floats from its right-hand neighbour at the same time, once with a
library copy (cudaMemcpy peer-to-peer through torch) and once with this repo's streaming kernel (mix_n reading a
peer-mapped source); the better of the two is "the link".  Fractions against the nominal 900 GB/s are printed beside it.
For a kernel that also computes, the roofline is the slower of its link time and its measured compute-only time."""
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

NOMINAL = 900e9


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
    # ---- link bandwidth, measured: all ranks pull 4*D bytes from their right neighbour simultaneously --------------------------
    from fedb200.parallel.peer_mem import _RawCuda
    nxt = (rank + 1) % G
    peer_acc = torch.as_tensor(_RawCuda(comm.peer_ptr(acc, nxt), D * 4), device=dev).view(torch.float32)
    ms_lib = timed(lambda: out.copy_(peer_acc), dev)
    ms_ker = timed(lambda: cuda_ops.mix_n([comm.peer_ptr(acc, nxt)], [1.0], out), dev)
    link_lib, link_ker = D * 4 / ms_lib * 1e3, D * 4 / ms_ker * 1e3
    LINK = max(link_lib, link_ker) if G > 1 else 770e9
    # ---- K1 fused FedAvg reduce (two-shot, + scale + bf16 shadow + broadcast) vs NCCL all_reduce + ATen --------------
    ms = timed(lambda: comm.weighted_mean(acc, 1.0, out, shadow), dev)
    bytes_dir = 2 * (G - 1) / G * D * 4 + (G - 1) / G * D * 2        # reduce-scatter reads + all-gather pushes (fp32 + bf16 shadow)
    rows.append(("fedavg_reduce (fp32 in, fp32 + bf16 out)", ms, bytes_dir / LINK * 1e3))
    ms = timed(lambda: comm.weighted_mean(acc, 1.0, out), dev)
    bytes_dir = 2 * (G - 1) / G * D * 4
    rows.append(("fedavg_reduce (fp32 in/out)", ms, bytes_dir / LINK * 1e3))
    ms = timed(lambda: comm.weighted_mean(acc, 1.0, out, max_ctas=32), dev)
    rows.append(("fedavg_reduce (fp32 in/out), 32-CTA grid (the form overlapped with compute)", ms, bytes_dir / LINK * 1e3))
    if comm.nvls_available():
        nv_a, nv_o = comm.alloc("b_nv_acc", D, multicast=True), comm.alloc("b_nv_out", D, multicast=True)
        nv_a.copy_(acc)
        ms = timed(lambda: comm.weighted_mean(nv_a, 1.0, nv_o, nvls=True), dev)
        # switch-reduced: a GPU sends its replica of every slice once (D) and receives its reduced slice + everyone's results (D/G + D)
        rows.append(("fedavg_reduce NVLS (multimem.ld_reduce / multimem.st, fp32 in/out)", ms, (1 + 1 / G) * D * 4 / LINK * 1e3))
    small_a, small_o = comm.alloc("b_small_a", 4096), comm.alloc("b_small_o", 4096)
    def small50():                      # 50 back-to-back calls: a single call's event time is dominated by the launch skew between ranks
        for _ in range(50):
            comm.weighted_mean(small_a, 1.0, small_o)
    ms = timed(small50, dev, iters=5) / 50
    rows.append(("one-shot all-reduce, 16 KB, per call over 50 back-to-back (latency-bound; roofline ~ 2 barriers x NVLink round trip)", ms, 0.004))
    buf = acc.clone()
    def nccl():
        dist.all_reduce(buf); buf.mul_(1.0 / G); shadow.copy_(buf)
    rows.append(("NCCL all_reduce + mul + bf16 cast (baseline)", timed(nccl, dev), bytes_dir / LINK * 1e3))
    # ---- K4 all-gather-fused Gram (tcgen05 + TMA over peer memory) vs all_gather + matmul -----------------------------
    deltas = comm.delta_buffer(1, D, torch.bfloat16)
    deltas[:, :D].copy_((torch.randn(1, D, device=dev) * 0.01).bfloat16())
    ms = timed(lambda: comm.gram(deltas, G), dev)
    def gram10():
        for _ in range(10):
            comm.gram(deltas, G)
    ms10 = timed(gram10, dev, iters=5) / 10
    remote = G * D * (G - 1) / (G * G) * 2                               # remote bytes read per GPU
    local_hbm = G * D / G * 2 / 6555e9 * 1e3
    rows.append((f"cosine_gram {G} peers (bf16, all-gather fused, single launch incl. N x N reduce)", ms, max(remote / LINK * 1e3, local_hbm)))
    rows.append((f"cosine_gram {G} peers, per call over 10 back-to-back (no inter-rank launch skew)", ms10, max(remote / LINK * 1e3, local_hbm)))
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
    dst_ptr = comm.peer_ptr(out, nxt)
    ms_hash = timed(lambda: cuda_ops.sha256_leaves(acc, None, 4096, leaves), dev)
    rows.append(("sha256 leaves only (no copy): the INT-pipe time of the hash", ms_hash, D * 4 * 2 / 6555e9 * 1e3))
    ms = timed(lambda: cuda_ops.sha256_leaves(acc, None, 4096, leaves, dst_ptr=dst_ptr), dev)
    rows.append(("sha256_send to peer (433 MB hashed while copied over NVLink); roofline = max(link, hash-only)", ms,
                 max(D * 4 / LINK * 1e3, ms_hash)))
    comm.check_errors()
    if rank == 0:
        res = {"n_gpus": G, "model": a.model, "params": D, "link_GBps_per_dir": round(LINK / 1e9, 1),
               "link_measured": {"library_peer_copy_GBps": round(link_lib / 1e9, 1), "kernel_peer_read_GBps": round(link_ker / 1e9, 1),
                                 "nominal_GBps": NOMINAL / 1e9, "how": "every rank pulls 4*D bytes from rank+1 at the same time; median of 10, max over ranks"},
               "rows": [{"kernel": n, "ms": round(t, 4), "roofline_ms": round(r, 4), "fraction_of_roofline": round(r / t, 3)} for n, t, r in rows]}
        print(json.dumps(res))
        print(f"\nlink measured: library peer copy {link_lib/1e9:.0f} GB/s, kernel peer read {link_ker/1e9:.0f} GB/s per direction "
              f"(nominal 900); roofline uses {LINK/1e9:.0f} GB/s")
        print(f"\n| kernel ({G} x B200, D = {D/1e6:.1f} M) | time ms | roofline ms | fraction of measured | of nominal 900 |\n|---|---|---|---|---|")
        for n, t, r in rows:
            print(f"| {n} | {t:.3f} | {r:.3f} | {r / t:.2f} | {r * LINK / NOMINAL / t:.2f} |")
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
