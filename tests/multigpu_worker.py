"""Worker run under torchrun (one process per GPU) by tests/test_multigpu.py: fused peer-memory collectives vs
torch.distributed NCCL + torch ops."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from fedb200.parallel import init_distributed, make_comm


def main():
    rank, world, local = init_distributed(torch.device("cuda"))
    dev = torch.device("cuda", local)
    comm = make_comm("peer", dev, rank, world)
    assert comm.name == "peer"
    torch.manual_seed(100 + rank)

    # ---- K1 fused FedAvg reduce: weighted mean with per-rank weights, odd length (scalar tail) ----------
    n = 3_000_003
    acc = comm.alloc("t_acc", n)
    out = comm.alloc("t_out", n)
    shadow = comm.alloc("t_shadow", n, torch.bfloat16)
    w = float(rank + 1)
    x = torch.randn(n, device=dev)
    acc.copy_(x * w)                                        # acc_r = n_r * w_r
    comm.weighted_mean(acc, w, out, shadow)
    ref = (x * w).clone()
    dist.all_reduce(ref)
    ref /= sum(range(1, world + 1))
    torch.cuda.synchronize()
    comm.check_errors()
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(shadow.float(), ref.bfloat16().float(), rtol=0, atol=4e-2)
    for it in range(3):                                      # repeated epochs reuse the flag words
        acc.copy_(x * (it + 1))
        comm.weighted_mean(acc, 1.0, out)
        ref = (x * (it + 1)).clone()
        dist.all_reduce(ref)
        ref /= world
        torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-6)

    # small grid (the form that overlaps compute on a side stream): same result, no co-residency requirement
    acc.copy_(x * w)
    comm.weighted_mean(acc, w, out, shadow, max_ctas=8)
    ref = (x * w).clone()
    dist.all_reduce(ref)
    ref /= sum(range(1, world + 1))
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-6)
    # every contributor masked out (weight 0 on every rank): the previous global model is kept
    keep = out.clone()
    acc.zero_()
    comm.weighted_mean(acc, 0.0, out)
    torch.testing.assert_close(out, keep, rtol=0, atol=0)
    # one-shot variant for small payloads (odd length -> scalar tail), interleaved with two-shot epochs
    ns = 4099
    sa, so = comm.alloc("t_small_acc", ns), comm.alloc("t_small_out", ns)
    for it in range(3):
        xs = torch.randn(ns, device=dev)
        sa.copy_(xs)
        comm.weighted_mean(sa, 1.0, so)                     # auto: <= 64 KB -> one-shot
        rs = xs.clone()
        dist.all_reduce(rs)
        rs /= world
        torch.testing.assert_close(so, rs, rtol=1e-5, atol=1e-6)
        comm.weighted_mean(acc, 1.0, out)                   # a two-shot call in between must not disturb the epochs
    # NVLS form (reduction inside the NVSwitch via multimem.ld_reduce / multimem.st) where the box exposes multicast
    if comm.nvls_available():
        na, no = comm.alloc("t_nv_acc", n, multicast=True), comm.alloc("t_nv_out", n, multicast=True)
        for it in range(2):
            na.copy_(x * (w + it))
            torch.cuda.synchronize(); dist.barrier()
            comm.weighted_mean(na, w + it, no, nvls=True)
            ref = (x * (w + it)).clone()
            dist.all_reduce(ref)
            ref /= sum(r + 1 + it for r in range(world))
            torch.testing.assert_close(no, ref, rtol=1e-5, atol=1e-6)
        keep_nv = no.clone()
        comm.weighted_mean(na, 0.0, no, nvls=True)            # all masked: untouched
        torch.testing.assert_close(no, keep_nv, rtol=0, atol=0)
        comm.check_errors()
    # slot exchange: device-side all-gather of small per-rank records
    tab = comm.alloc("t_tab", world * 48, torch.int32).view(world, 48)
    for it in range(2):
        tab[rank].copy_(torch.arange(48, device=dev, dtype=torch.int32) + 1000 * rank + it)
        comm.exchange_slots(tab)
        want = torch.stack([torch.arange(48, dtype=torch.int32) + 1000 * r + it for r in range(world)]).to(dev)
        torch.cuda.synchronize()
        assert torch.equal(tab, want), "slot exchange mismatch"
        dist.barrier()                                      # nobody rewrites its row while a peer still compares
    comm.check_errors()

    # ---- K4 all-gather-fused Gram over peer memory ---------------------------------------------------------
    P, D = 3, 64 * 997 + 8
    deltas = comm.alloc("t_deltas", P * D, torch.bfloat16).view(P, D)
    deltas.copy_((torch.randn(P, D, device=dev) * 0.1).bfloat16())
    g = comm.gram(deltas, world * P)
    allrows = [torch.empty_like(deltas) for _ in range(world)]
    dist.all_gather(allrows, deltas.contiguous())
    full = torch.cat(allrows, 0).float()
    gref = full @ full.T
    torch.cuda.synchronize()
    comm.check_errors()
    err = (g - gref).abs().max().item()
    assert err < 2e-3 * gref.abs().max().item() + 1e-3, f"gram err {err}"

    # same through the segment-stacked layout the runtime uses for few clients
    dseg = comm.delta_buffer(P, D - 8, torch.bfloat16)
    dseg[:, : D - 8].copy_((torch.randn(P, D - 8, device=dev) * 0.1).bfloat16())
    g2 = comm.gram(dseg, world * P)
    allseg = [torch.empty_like(dseg) for _ in range(world)]
    dist.all_gather(allseg, dseg.contiguous())
    f2 = torch.cat(allseg, 0).float()
    torch.cuda.synchronize()
    comm.check_errors()
    err = (g2 - f2 @ f2.T).abs().max().item()
    assert err < 2e-3 * (f2 @ f2.T).abs().max().item() + 1e-3, f"segmented gram err {err}"

    # ---- K2 routed mixing: out[i] = sum_j W[i, j] slot_j over peer memory ----------------------------------------
    N, Dm = world * 2, 500_000
    slots = comm.alloc("t_slots", 2 * Dm).view(2, Dm)
    outm = comm.alloc("t_mix", 2 * Dm).view(2, Dm)
    slots.copy_(torch.randn(2, Dm, device=dev))
    W = np.random.default_rng(0).random((N, N))
    W[W < 0.3] = 0.0
    W /= W.sum(1, keepdims=True)
    comm.mix_rows(slots, W[rank * 2: rank * 2 + 2], outm, N)
    alls = [torch.empty_like(slots) for _ in range(world)]
    dist.all_gather(alls, slots.contiguous())
    S = torch.cat(alls, 0)
    refm = torch.tensor(W[rank * 2: rank * 2 + 2], dtype=torch.float32, device=dev) @ S
    torch.cuda.synchronize()
    comm.check_errors()
    torch.testing.assert_close(outm, refm, rtol=1e-4, atol=1e-5)

    # same exchange again (second epoch of the fused kernel's flag words) and, with >= 3 GPUs, two-hop routes executed
    # through a relay GPU: the result must not depend on how the bytes travelled
    from fedb200.p2p import RelayPlan
    from fedb200.parallel import Placement
    comm.placement = Placement.block(N, world)
    slots.copy_(torch.randn(2, Dm, device=dev))
    relays = {}
    if world >= 3:
        for i in range(N):
            for j in range(N):
                if W[i, j] != 0.0 and i != j and (i + j) % 3 == 0:
                    relays[(i, j)] = (max(i, j) + 2) % N           # some third peer
    plan = RelayPlan(W, comm.placement.owner, relays, comm.relay_slots)
    comm.mix_rows(slots, W[rank * 2: rank * 2 + 2], outm, N, plan=plan if plan.any else None)
    dist.all_gather(alls, slots.contiguous())
    refm = torch.tensor(W[rank * 2: rank * 2 + 2], dtype=torch.float32, device=dev) @ torch.cat(alls, 0)
    torch.cuda.synchronize()
    comm.check_errors()
    torch.testing.assert_close(outm, refm, rtol=1e-4, atol=1e-5)
    n_relayed = len(plan.via)
    assert world < 3 or n_relayed > 0, "the relay path was not exercised"
    comm.placement = None

    # ---- whole serverless round through the runtime on the fused backend ------------------------------------------
    from fedb200.config import FedConfig, ModelConfig
    from fedb200.runtime import FedRuntime
    cfg = FedConfig(mode="serverless", num_clients=world, num_rounds=2, device="cuda", backend="fused", comm="peer")
    cfg.model = ModelConfig.preset("bert-mini", num_labels=2)
    cfg.data.population, cfg.data.train_per_client, cfg.data.test_per_client = 800, 32, 32
    cfg.data.max_seq_len, cfg.data.batch_size, cfg.ledger = 128, 16, True
    cfg.anomaly_detector = "cosine_pagerank"
    rt = FedRuntime(cfg, echo=False, comm=comm)               # reuse the already-bootstrapped backend
    res = rt.run()
    torch.cuda.synchronize()
    comm.check_errors()
    mine = rt.slots[0].clone()
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    keep = [i for i in range(world) if i not in rt.masked]
    torch.testing.assert_close(rt.global_flat, torch.stack([gathered[i] for i in keep]).mean(0), rtol=1e-5, atol=1e-6)
    chk = torch.tensor([float(rt.global_flat.double().sum())], device=dev, dtype=torch.float64)
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert lo.item() == hi.item(), "ranks disagree on the global model"
    assert rt.ledger.height == 2 * world and rt.ledger.validate() == -1
    heads = [None] * world
    dist.all_gather_object(heads, rt.ledger.head.hex())
    assert len(set(heads)) == 1, "ledgers diverged across ranks"
    # ---- same with a static mask: the aggregate overlaps the local evaluation on the side stream, 2 peers per GPU --------
    cfg1 = FedConfig(mode="serverless", num_clients=2 * world, num_rounds=2, device="cuda", backend="fused", comm="peer")
    cfg1.model = ModelConfig.preset("bert-mini", num_labels=2)
    cfg1.data.population, cfg1.data.train_per_client, cfg1.data.test_per_client = 800, 32, 32
    cfg1.data.max_seq_len, cfg1.data.batch_size, cfg1.ledger = 128, 16, True
    rt1 = FedRuntime(cfg1, echo=False, comm=comm)
    res1 = rt1.run()
    torch.cuda.synchronize()
    comm.check_errors()
    mine = rt1.slots[: len(rt1.local_ids)].sum(0)
    dist.all_reduce(mine)
    torch.testing.assert_close(rt1.global_flat, mine / (2 * world), rtol=1e-5, atol=1e-6)
    assert rt1.ledger.height == 2 * 2 * world and rt1.ledger.validate() == -1
    heads = [None] * world
    dist.all_gather_object(heads, (rt1.ledger.head.hex(), res1["last"]["local_accuracy"]))
    assert len(set(h[0] for h in heads)) == 1 and all(h[1] == heads[0][1] for h in heads), "records diverged across ranks"
    assert len(heads[0][1]) == 2 * world
    # ---- the same round with the switch-side (NVLS) aggregate: --nvls allocates slots / global model on a multicast object ------
    if comm.nvls_available():
        import copy
        cfgn = copy.deepcopy(cfg1)
        cfgn.nvls, cfgn.num_clients = True, world
        rtn = FedRuntime(cfgn, echo=False, comm=comm)
        assert rtn.nvls
        rtn.run()
        torch.cuda.synchronize()
        comm.check_errors()
        mine = rtn.slots[0].clone()
        dist.all_reduce(mine)
        torch.testing.assert_close(rtn.global_flat, mine / world, rtol=1e-5, atol=1e-6)
        assert rtn.ledger.validate() == -1
    # ---- asynchronous gossip across GPUs: schedule-driven executor, device-side pub/done flags ---------------------
    cfg2 = FedConfig(mode="serverless", num_clients=2 * world, num_rounds=3, device="cuda", backend="fused", comm="peer",
                     asynchronous=True, ledger=True, topology="full")
    cfg2.model = ModelConfig.preset("bert-mini", num_labels=2)
    cfg2.data.population, cfg2.data.train_per_client, cfg2.data.test_per_client = 800, 32, 32
    cfg2.data.max_seq_len, cfg2.data.batch_size = 128, 16
    cfg2.async_epoch_s, cfg2.async_spread = 0.001, 1.0          # short modelled epochs so versions overlap and slots recycle
    cfg2.async_fanin, cfg2.async_chunk_kb = 3, 256            # several landing flags per publish even for the small model
    rt2 = FedRuntime(cfg2, echo=False, comm=comm)
    rt2._audit_async = True
    ares = rt2.run()["last"]
    torch.cuda.synchronize()
    comm.check_errors()
    assert ares["async_mode"] == "observed" and ares["torn_chunks"] == 0
    assert ares["mixes"] == 3 * 2 * world and rt2.ledger.height == ares["mixes"] and rt2.ledger.validate() == -1
    chk = torch.tensor([float(rt2.global_flat.double().sum())], device=dev, dtype=torch.float64)
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert lo.item() == hi.item(), "ranks disagree on the async consensus model"
    heads = [None] * world
    dist.all_gather_object(heads, rt2.ledger.head.hex())
    assert len(set(heads)) == 1, "async ledgers diverged across ranks"
    # the consensus equals the mean of every peer's final state
    mine = rt2.async_state[: len(rt2.local_ids)].sum(0)
    dist.all_reduce(mine)
    torch.testing.assert_close(rt2.global_flat, mine / (2 * world), rtol=1e-5, atol=1e-6)
    # replay: every mix equals the host oracle evaluated on the versions the device kernel OBSERVED and logged
    from fedb200.p2p.observed import replay_mix
    pubs = [None] * world
    dist.all_gather_object(pubs, rt2.async_audit["published"])
    published = {k: v for d in pubs for k, v in d.items()}
    cross = 0
    for r in rt2.async_records:
        want = replay_mix(rt2.async_audit["trained"][(r["peer"], r["epoch"])], r["versions"], published, r["sources"],
                          r["epoch"], rt2.async_table)
        torch.testing.assert_close(rt2.async_audit["published"][(r["peer"], r["epoch"])], want, rtol=1e-5, atol=1e-6)
        cross += sum(1 for j, v in zip(r["sources"], r["versions"]) if v > 0 and j not in rt2.local_ids)
    tot = torch.tensor([float(cross)], device=dev)
    dist.all_reduce(tot)
    assert tot.item() > 0, "no mix consumed a version published on another GPU"
    # on-line scoring in asynchronous mode: fused Gram + device PageRank every epoch, mask consumed on the device
    cfg4 = FedConfig(mode="serverless", num_clients=2 * world, num_rounds=3, device="cuda", backend="fused", comm="peer",
                     asynchronous=True, topology="full", anomaly_detector="cosine_pagerank", inject_anomalies=1,
                     inject_kind="signflip", async_score_every=1)
    cfg4.model = ModelConfig.preset("bert-mini", num_labels=2)
    cfg4.data.population, cfg4.data.train_per_client, cfg4.data.test_per_client = 800, 32, 32
    cfg4.data.max_seq_len, cfg4.data.batch_size = 128, 16
    rt4 = FedRuntime(cfg4, echo=False, comm=comm)
    a4 = rt4.run()["last"]
    comm.check_errors()
    allm = [None] * world
    dist.all_gather_object(allm, a4["masked"])
    assert all(m == allm[0] for m in allm), "ranks disagree on the anomaly mask"
    # ---- ring gossip through the runtime: per-peer states, fused mix kernel, executed routes in the ledger ----------------
    cfg3 = FedConfig(mode="serverless", num_clients=2 * world, num_rounds=2, device="cuda", backend="fused", comm="peer",
                     topology="reference10" if 2 * world == 10 else "ring", ledger=True)
    cfg3.model = ModelConfig.preset("bert-mini", num_labels=2)
    cfg3.data.population, cfg3.data.train_per_client, cfg3.data.test_per_client = 800, 32, 32
    cfg3.data.max_seq_len, cfg3.data.batch_size = 128, 16
    rt3 = FedRuntime(cfg3, echo=False, comm=comm)
    rt3.run()
    torch.cuda.synchronize()
    comm.check_errors()
    from fedb200.p2p import apply_mask, mixing_matrix
    Wr = apply_mask(mixing_matrix(rt3.topo.adjacency(), "metropolis"), rt3.masked)
    alls3 = [torch.empty_like(rt3.slots) for _ in range(world)]
    dist.all_gather(alls3, rt3.slots.contiguous())
    cat3 = torch.cat(alls3, 0)                                # (rank, slot) order -> peer-id order
    S3 = torch.stack([cat3[rt3.placement.owner(k)[0] * rt3.P + rt3.placement.owner(k)[1]] for k in range(2 * world)])
    want3 = torch.tensor(Wr[rt3.local_ids], dtype=torch.float32, device=dev) @ S3
    torch.testing.assert_close(rt3.peer_state[: len(rt3.local_ids)], want3, rtol=1e-4, atol=1e-5)
    assert rt3.ledger.validate() == -1 and rt3.ledger.height == 2 * 2 * world
    heads = [None] * world
    dist.all_gather_object(heads, rt3.ledger.head.hex())
    assert len(set(heads)) == 1, "gossip ledgers diverged across ranks"
    if rank == 0:
        print("MULTIGPU_OK", world, res["global_accuracies"], "async mixes", ares["mixes"], "mean staleness",
              round(ares["mean_staleness"], 2), "async masked", a4["masked"], "targets", rt4.injector.targets())
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
