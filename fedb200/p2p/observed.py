"""Asynchronous gossip from OBSERVED versions (K3 host side; the multi-GPU executor behind ``--async``).

The reference specifies asynchrony only as a timing formula (an asynchronous exchange costs the max, not the sum, of the
per-destination transfer times, Medical_Transcriptions_All_graphs.ipynb:979-980, 997-999).  Round 1 executed a
virtual-time schedule built from MODELLED epoch durations, so a peer that was slow in reality still blocked its consumers
in schedule order.  Here nothing is scheduled:

* every peer trains, mixes and publishes at its own pace; version ``e`` goes to slot ``e % R`` of its outbox ring, its
  ``pub`` word is bumped first and one landing flag per chunk is released as the publish kernel writes the slot;
* a mix (``gossip_mix_async_kernel``) reads its sources' ``pub`` words on the device — whatever they hold at that moment —
  fixes those versions, discounts each by its staleness ``alpha_j = (mix_rate / n_live) / (1 + tau_j)^decay`` and walks the
  chunks as they land; no stream ever waits for a peer to reach a particular epoch;
* the chosen versions are logged per mix, so the host can replay every mix exactly (``replay_mix``); fan-in is bounded
  (``--async-fanin``, default 4 nearest in-neighbours) so a BERT-large mix reads 4 remote models, not 63;
* with ``--anomaly-detector cosine_pagerank --async-score-every K`` the fused Gram + on-device PageRank re-score every K
  epochs and write a device mask that the mix kernel consults — the mask never visits the host inside the loop.
"""
from __future__ import annotations

import time
from typing import Dict, List

import numpy as np
import torch


def stale_table(mix_rate: float, decay: float) -> List[float]:
    """``mix_rate / (1 + tau)^decay`` for tau = 0..31 as float32 values (the device divides by the live-source count)."""
    return [float(np.float32(mix_rate / float((1 + t) ** decay))) for t in range(32)]


def pick_sources(topo, nbytes: float, peer: int, fanin: int, excluded) -> List[int]:
    """The ``fanin`` in-neighbours of ``peer`` with the fastest channel towards it (hop time = size / bandwidth)."""
    adj, hop = topo.adjacency(), topo.hop_seconds(nbytes)
    cand = [j for j in range(topo.n) if j != peer and adj[peer, j] > 0 and j not in excluded]
    cand.sort(key=lambda j: (hop[j][peer], j))
    return cand[: max(int(fanin), 0)]


def replay_mix(trained: torch.Tensor, versions: List[int], published: Dict, sources: List[int], my_epoch: int,
               table: List[float]) -> torch.Tensor:
    """Host oracle of one mix from its logged versions: ``published[(peer, version)]`` are the published states."""
    live = [(j, v) for j, v in zip(sources, versions) if v != 0]
    acc = torch.zeros_like(trained, dtype=torch.float32)
    tot = np.float32(0.0)
    ws = []
    for j, v in live:
        tau = min(max(my_epoch - v, 0), 31)
        w = np.float32(table[tau]) / np.float32(len(live))
        ws.append(w)
        tot = np.float32(tot + w)
    acc.add_(trained.float(), alpha=float(np.float32(1.0) - tot))
    for (j, v), w in zip(live, ws):
        acc.add_(published[(j, v)].float(), alpha=float(w))
    return acc


def run_async_observed(rt) -> Dict:
    """Run ``cfg.num_rounds`` local epochs per peer with observed-version gossip; returns the run summary."""
    from ..data.build import client_loaders, global_test_loader
    from ..ledger.block import BROADCAST
    from ..parallel.placement import owner_of
    from .path_opt import info_passing_time
    cfg, N, D, comm, P, dev = rt.cfg, rt.N, rt.D, rt.comm, rt.P, rt.device
    on_gpu = dev.type == "cuda"
    if rt.world > 1 and comm.name != "peer":
        raise NotImplementedError("multi-rank asynchronous gossip needs the peer-memory backend (device flags)")
    E, R = cfg.num_rounds, max(int(cfg.async_ring), 2)
    chunk_elems = max(4, (cfg.async_chunk_kb * 1024 // 4) // 4 * 4)
    n_chunks = (D + chunk_elems - 1) // chunk_elems
    static_masked = set(rt.static_masked)
    table = stale_table(0.5, cfg.staleness_decay)
    F = min(int(cfg.async_fanin), 8)
    outbox = comm.alloc("obs_outbox", P * R * D).view(P, R, D)
    cflags = comm.alloc("obs_cflags", P * R * n_chunks, torch.int32).view(P, R, n_chunks)
    pub = comm.alloc("obs_pub", max(P, 4), torch.int32)
    state = torch.empty(P, D, dtype=torch.float32, device=dev)
    decision = torch.zeros(P, 16, dtype=torch.int32, device=dev)
    log = torch.zeros(E * P + 1, F + 1, dtype=torch.int32, device=dev)
    err = torch.zeros(4, dtype=torch.int32, device=dev)
    mask_dev = torch.zeros(N, dtype=torch.int32, device=dev)
    for k in static_masked:
        mask_dev[k] = 1
    init_flat = rt.global_flat.clone()
    for li in range(len(rt.local_ids)):
        state[li].copy_(rt.global_flat)
    sources = {k: pick_sources(rt.topo, rt.nbytes, k, F, static_masked) for k in rt.local_ids}
    ledger_on = rt.ledger is not None
    root_table = torch.zeros(E * P + 1, 32, dtype=torch.uint8, device=dev) if ledger_on else None
    if ledger_on and on_gpu:
        from ..ops import cuda_ops
        nch = (rt.nbytes + rt.ledger.chunk_bytes - 1) // rt.ledger.chunk_bytes
        leaves_ws = torch.empty(nch, 32, dtype=torch.uint8, device=dev)
        scratch_ws = torch.empty((nch + 1) // 2 + 1, 32, dtype=torch.uint8, device=dev)
    audit = getattr(rt, "_audit_async", False)
    rt.async_audit = {"trained": {}, "published": {}, "mixed": {}}
    if on_gpu:
        torch.cuda.synchronize()
    comm.barrier()
    events, mix_events = [], []
    samples = tokens = 0
    launch_id = [0] * P
    t0 = time.perf_counter()
    for e in range(1, E + 1):
        for li, k in enumerate(rt.local_ids):
            if k in static_masked:
                continue
            tr, _ = client_loaders(rt.corpus, cfg.data, k, e - 1, pin=on_gpu)
            with rt.timer.phase("train"):
                rt.runner.load_flat(state[li])
                rt.runner.sync_shadow()
                rt.runner.train_epochs(tr, cfg.local_epochs, lazy=True)
                if rt.injector.maybe_corrupt(k, rt.runner.flat, state[li], e - 1) and hasattr(rt.runner, "mark_dirty"):
                    rt.runner.mark_dirty()
            samples, tokens = samples + tr.num_samples, tokens + tr.num_tokens()
            row = (e - 1) * P + li
            srcs = sources[k]
            if audit:
                rt.async_audit["trained"][(k, e)] = rt.runner.flat.detach().clone().cpu()
            with rt.timer.phase("aggregate"):
                launch_id[li] += 1
                if on_gpu:
                    from ..ops import cuda_ops
                    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    ev0.record()
                    rings, flags, pubs = [], [], []
                    for j in srcs:
                        rj, lj = owner_of(j, N, rt.world)
                        rings.append(comm.peer_ptr(outbox[lj], rj))
                        flags.append(comm.peer_ptr(cflags[lj], rj))
                        pubs.append(comm.peer_ptr(pub, rj) + 4 * lj)
                    cuda_ops.gossip_mix_async(rt.runner.flat, state[li], rings, flags, pubs, srcs, mask_dev, decision[li],
                                              log[row], err.data_ptr(), R, chunk_elems, e, launch_id[li], table,
                                              timeout_ms=getattr(comm, "timeout_ms", 20000))
                    comm.launches += 1
                else:                                  # CPU plumbing twin (single process): same rule, torch ops
                    vers = [0 if int(mask_dev[j]) else int(pub[owner_of(j, N, rt.world)[1]]) for j in srcs]
                    published = {(j, v): outbox[owner_of(j, N, rt.world)[1], v % R] for j, v in zip(srcs, vers) if v}
                    state[li].copy_(replay_mix(rt.runner.flat, vers, published, srcs, e, table))
                    log[row, : len(srcs)] = torch.tensor(vers, dtype=torch.int32) if srcs else log[row, :0]
            with rt.timer.phase("publish"):
                slot = outbox[li, e % R]
                comm.flag_set(pub, li, e)                                   # publish STARTED: consumers may pick version e
                if on_gpu:
                    if e > R:
                        # the slot still holds version e - R with landing flags that say so: retire the flags BEFORE its bytes
                        # change, so a reader that is still on that version finds the mismatch in its post-read check (without
                        # this a read lapped in the middle of a chunk would go unnoticed: utils/protocol_check.py::check_ring)
                        cuda_ops.set_flags(cflags[li, e % R], 0, n_chunks)
                        comm.launches += 1
                    if ledger_on:
                        cuda_ops.sha256_leaves(state[li], slot, rt.ledger.chunk_bytes, leaves_ws)
                        cuda_ops.merkle_root(leaves_ws, scratch_ws, root_table[row])
                        cuda_ops.set_flags(cflags[li, e % R], e, n_chunks)
                        comm.launches += 3
                    else:
                        cuda_ops.copy_flagged(state[li], slot, chunk_elems, cflags[li, e % R], e)
                        comm.launches += 1
                    ev1.record()
                    mix_events.append((ev0, ev1))
                else:
                    slot.copy_(state[li])
                    cflags[li, e % R].fill_(e)
                    if ledger_on:
                        from ..ledger.merkle import merkle_root
                        root_table[row] = torch.frombuffer(bytearray(merkle_root(slot.numpy().tobytes(), rt.ledger.chunk_bytes)),
                                                           dtype=torch.uint8)
            if audit:
                rt.async_audit["published"][(k, e)] = state[li].detach().clone().cpu()
            events.append((row, k, e, list(srcs)))
        if cfg.anomaly_detector == "cosine_pagerank" and cfg.async_score_every > 0 and e % cfg.async_score_every == 0 and e < E:
            _score_on_device(rt, state, init_flat, mask_dev)
    if on_gpu:
        torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    comm.barrier()
    comm.check_errors()
    if int(err[0]) != 0:
        raise RuntimeError("asynchronous mix timed out waiting for a chunk flag (a publish never landed)")
    return _finish_observed(rt, state, mask_dev, log, root_table, events, mix_events, samples, tokens, wall, R, table)


def _score_on_device(rt, state: torch.Tensor, init_flat: torch.Tensor, mask_dev: torch.Tensor) -> None:
    """Cosine / PageRank scoring of the peers' accumulated updates (state - initial model); the resulting mask stays on the
    device and is what the following mixes consult.  The fused Gram kernel contains a cross-rank barrier, so this is the one
    point where the asynchronous ranks meet (every ``--async-score-every`` epochs)."""
    cfg, N, P, D, comm = rt.cfg, rt.N, rt.P, rt.D, rt.comm
    on_gpu = rt.device.type == "cuda"
    with rt.timer.phase("score"):
        if rt.deltas is None:
            rt.deltas = comm.delta_buffer(P, D, torch.bfloat16 if on_gpu else torch.float32)
        for li in range(len(rt.local_ids)):
            if on_gpu:
                from ..ops import cuda_ops
                cuda_ops.delta_bf16(state[li], init_flat, rt.deltas[li][:D])
            else:
                torch.sub(state[li], init_flat, out=rt.deltas[li][:D])
        g = comm.gram(rt.deltas, N)[:N, :N].contiguous()
        if on_gpu:
            from ..ops import cuda_ops
            ranks, mask, _ = cuda_ops.pagerank(g, mode=1, low_only=True)
            mean = ranks.mean()
            keep_guard = (ranks - mean).abs() > 0.1 * mean           # same guard as anomaly.cosine.score_updates
            mask_dev.copy_(torch.maximum(mask_dev, (mask.to(torch.bool) & keep_guard).to(torch.int32)))
        else:
            from ..anomaly import score_updates
            bad, _, _ = score_updates(g.to(torch.float64).numpy())
            for k in bad:
                mask_dev[k] = 1


def _finish_observed(rt, state, mask_dev, log, root_table, events, mix_events, samples, tokens, wall, R, table) -> Dict:
    from ..data.build import global_test_loader
    from ..ledger.block import BROADCAST
    from .path_opt import info_passing_time
    cfg, N, comm = rt.cfg, rt.N, rt.comm
    on_gpu = rt.device.type == "cuda"
    masked = [int(i) for i in torch.nonzero(mask_dev).flatten().cpu().tolist()]
    keep_w = [0.0 if k in masked else 1.0 for k in rt.local_ids]
    L = len(rt.local_ids)
    rt.acc.zero_()
    for i in range(L):
        if keep_w[i]:
            rt.acc.add_(state[i])
    comm.weighted_mean(rt.acc, float(sum(keep_w)), rt.global_flat)
    rt.runner.load_flat(rt.global_flat)
    rt.runner.sync_shadow()
    gl, ga = rt.runner.evaluate(global_test_loader(rt.corpus, cfg.data, cfg.num_rounds, pin=on_gpu))
    rt.global_accuracies.append(float(ga))
    logs = log.cpu().numpy()
    F = logs.shape[1] - 1
    records = [{"peer": k, "epoch": e, "sources": srcs, "versions": [int(v) for v in logs[row, : len(srcs)]],
                "torn": int(logs[row, len(srcs)])} for row, k, e, srcs in events]
    torn = sum(r["torn"] for r in records)
    stal = [e - v for r in records for e, v in [(r["epoch"], x) for x in r["versions"]] if v > 0]
    if rt.ledger is not None:
        rows = root_table.cpu().numpy()
        local = [(e, k, bytes(rows[row]).hex(), srcs[:16]) for row, k, e, srcs in events]
        for e, k, root_hex, srcs in sorted(sum(comm.all_gather_object(local), [])):
            rt.ledger.append(e, k, BROADCAST, bytes.fromhex(root_hex), rt.nbytes, 0, srcs, timestamp_ns=e * 1_000_000 + k)
        rt.ledger.assert_valid()
    mix_ms = [a.elapsed_time(b) for a, b in mix_events] if mix_events else [0.0]
    tot = comm.reduce_scalars([samples, tokens, len(records), torn], "sum")
    mx = comm.reduce_scalars([wall, max(mix_ms), float(np.mean(mix_ms))], "max")
    model = info_passing_time(rt.topo, rt.nbytes, 0, removed=masked)
    res = {"accuracy": float(ga), "loss": float(gl), "mixes": int(tot[2]), "torn_chunks": int(tot[3]), "masked": masked,
           "wall_s": mx[0], "samples_per_s": tot[0] / mx[0] if mx[0] else None, "tokens_per_s": tot[1] / mx[0] if mx[0] else None,
           "async_mode": "observed", "fanin": F, "ring": R,
           # measured on the device: one peer's mix + publish (its whole information exchange of an epoch); async = max over
           # peers — beside the analytical model of the reference (size / bandwidth along routes)
           "info_passing_measured_ms": {"max_over_peers": mx[1], "mean_max_over_ranks": mx[2]},
           "info_passing": model,
           "mean_staleness": float(np.mean(stal)) if stal else 0.0, "max_staleness": int(max(stal)) if stal else 0}
    rt.async_records, rt.async_state, rt.async_table = records, state, table
    res["phases_ms"] = rt.timer.device_ms()
    rt.log.log("async_done", **res)
    return res
