"""Federated round loops: server (FedAvg) and serverless (gossip), sync and async.

What the reference does (call stacks in SURVEY §3):

* server mode — Flower ``Server.fit``: per round sample all clients, each ``fit``s from the global
  parameters, ``aggregate_fit`` = weighted mean, then every client ``evaluate``s the new global
  model and the metrics are example-weighted (src/Servercase/server_IID_IMDB.py:205-218);
* serverless mode — a Python ``for`` over clients that train one after another on ONE shared model
  object, snapshot their parameters, and an unweighted mean of the snapshots becomes the next
  global model; global accuracy on a fresh random test subset; ``save_pretrained`` every round
  (src/Serverlesscase/serverless_IID_IMDB.py:255-285).

Here clients/peers are spread over ranks (one process per GPU), weights stay on the device in flat
arenas, and aggregation / mixing / scoring / hashing are single fused kernels through ``Comm``.
``serverless_semantics="handoff"`` reproduces the reference's sequential chain for comparison;
the default ``"parallel"`` trains every peer from its own state concurrently (true gossip).
"""
from __future__ import annotations

import os
import time
from dataclasses import asdict
from typing import Dict, List, Optional

import numpy as np
import torch

from .anomaly import AnomalyInjector, detect, score_updates
from .config import FedConfig
from .data.build import build_corpus, client_loaders, global_test_loader
from .ledger import Ledger
from .ledger.block import BROADCAST
from .models.hf_io import load_state, save_pretrained
from .p2p import AsyncGossipEngine, RelayPlan, Topology, apply_mask, info_passing_time, mixing_matrix, optimise_route
from .parallel import init_distributed, local_peer_ids, make_comm, peers_per_rank
from .server.fedavg import FedAvg, weighted_average
from .trainer.local import make_runner, resolve_device
from .utils import MetricsLogger, OverheadMeter, PhaseTimer
from .utils.readback import HostReadback
from contextlib import nullcontext as _nullctx


class FedRuntime:
    def __init__(self, cfg: FedConfig, logger: Optional[MetricsLogger] = None, echo: bool = True, comm=None):
        """``comm``: an already bootstrapped communication backend to share (several runtimes in one process, tests);
        the runtime then does not close it."""
        self.cfg = cfg
        dev = resolve_device(cfg)
        self.rank, self.world, self.local_rank = init_distributed(dev)
        self.device = torch.device("cuda", self.local_rank) if dev.type == "cuda" else dev
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)
        self.overhead = OverheadMeter()                                   # before data/model load, like :59-63
        self.log = logger or MetricsLogger(cfg.metrics_path, echo=echo, rank=self.rank)
        self._own_comm = comm is None
        self.comm = comm if comm is not None else make_comm(cfg.comm, self.device, self.rank, self.world)
        self.N = cfg.num_clients
        self.P = peers_per_rank(self.N, self.world)
        cfg.data.max_seq_len = min(cfg.data.max_seq_len, cfg.model.max_position)
        self.corpus = build_corpus(cfg.data, cfg.model, self.N, cfg.seed, vocab_path=cfg.vocab_path or cfg.init_checkpoint)
        self.placement = self._make_placement()
        self.comm.placement = self.placement
        self.local_ids = self.placement.local_ids(self.rank)
        self.runner = make_runner(cfg, self.device, cfg.seed)
        self.D = self.runner.flat.numel()
        self.nbytes = self.D * 4
        self.nvls = bool(cfg.nvls) and self.world > 1 and self.comm.nvls_available()
        self.global_flat = self.comm.alloc("global", self.D, multicast=self.nvls)
        self.global_flat.copy_(self.runner.flat)
        self.slots = self.comm.alloc("slots", self.P * self.D, multicast=self.nvls).view(self.P, self.D)
        self.acc = self.comm.alloc("acc", self.D, multicast=self.nvls)
        self.deltas = None
        self.timer = PhaseTimer(self.device.type == "cuda")
        self.readback = HostReadback(self.device)
        self._loader_cache: Dict = {}
        self.topo = Topology.from_name(cfg.topology, self.N, cfg.seed)
        self.injector = AnomalyInjector(self.N, cfg.inject_anomalies, cfg.inject_kind, seed=cfg.seed)
        if cfg.inject_kind == "slowlink" and cfg.inject_anomalies > 0:
            # network-level fault: the targets' channels run at 5 % bandwidth; the graph detectors see it in the
            # topology (the reference's anomaly notion: a node whose channel bandwidths are off)
            self.topo = self.topo.slow_links(self.injector.targets(), 0.05)
            self.injector.count = 0                      # no weight corruption for this kind
        self.ledger = Ledger(chunk_bytes=4096) if cfg.ledger else None
        self.strategy = FedAvg(fraction_fit=1.0, fraction_evaluate=1.0,
                               evaluate_metrics_aggregation_fn=weighted_average, seed=cfg.seed)
        self.global_accuracies: List[float] = []
        self.masked: List[int] = []
        self.start_round = 0
        self.peer_state = None
        self.ckpt_writer = None
        from .utils.records import RoundRecords
        self.records = RoundRecords(self.comm, self.world, self.rank, self.P, self.device)
        self._route_cache: Dict = {}
        self._agg_pending = None
        self._agg_stream = torch.cuda.Stream(self.device, priority=-1) if self.device.type == "cuda" else None
        if cfg.mode == "serverless" and cfg.serverless_semantics == "parallel" and not self._full_uniform():
            self.peer_state = self.comm.alloc("state", self.P * self.D).view(self.P, self.D)
        if cfg.resume and cfg.checkpoint_dir and os.path.exists(os.path.join(cfg.checkpoint_dir, "config.json")):
            self._resume()
        if self.peer_state is not None and not getattr(self, "_peer_state_restored", False):
            for li in range(len(self.local_ids)):            # after _resume(): start from the checkpointed consensus
                self.peer_state[li].copy_(self.global_flat)
        static = detect(self.topo, cfg.anomaly_detector) if cfg.anomaly_detector not in ("none", "cosine_pagerank") else []
        self.static_masked = static
        if getattr(self, "_placement_loads", None):
            blk, bal = self._placement_loads
            self.log.log("placement", kind="balanced", max_tokens_block=max(blk), max_tokens_balanced=max(bal),
                         mean_tokens=sum(bal) / len(bal))
        self.log.log("init", mode=cfg.mode, clients=self.N, world=self.world, device=str(self.device),
                     backend=self.runner.backend, comm=self.comm.name, params=self.D, model=cfg.model.name,
                     static_masked=static)

    # ------------------------------------------------------------------------------------------
    def _make_placement(self):
        """Token-balanced (LPT) client -> GPU placement for the synchronous modes when clients outnumber GPUs; the cost of
        a client is the token count of its round-0 training set (static for the shard / Dirichlet splits, representative
        for IID resampling).  The asynchronous executors keep the block placement their schedules are written for."""
        from .parallel import Placement
        cfg = self.cfg
        if cfg.placement != "balanced" or self.world == 1 or self.N <= self.world or cfg.asynchronous:
            return Placement.block(self.N, self.world)
        cost = [client_loaders(self.corpus, cfg.data, k, 0)[0].num_tokens() for k in range(self.N)]
        pl = Placement.balanced(self.N, self.world, cost)
        self._placement_loads = (Placement.block(self.N, self.world).loads(cost), pl.loads(cost))
        return pl

    def _full_uniform(self) -> bool:
        return self.cfg.topology == "full" and not self.cfg.asynchronous

    def _resume(self) -> None:
        import json
        path = self.cfg.checkpoint_dir
        state = load_state(path)
        self.runner.load_state_dict(state)
        self.runner.sync_shadow()
        self.global_flat.copy_(self.runner.flat)
        fs = os.path.join(path, "fed_state.json")
        if os.path.exists(fs):
            with open(fs) as fh:
                meta = json.load(fh)
            self.start_round = int(meta.get("round", -1)) + 1
            self.global_accuracies = list(meta.get("global_accuracies", []))
            lp = os.path.join(path, "ledger.json")
            if self.ledger is not None and os.path.exists(lp):
                self.ledger = Ledger.load(lp)
                self.ledger.assert_valid()
        # gossip topologies keep one state per peer: each rank restores the states of its own peers
        pp = os.path.join(path, f"peer_states_rank{self.rank}.safetensors")
        if self.peer_state is not None and os.path.exists(pp):
            from safetensors.torch import load_file
            saved = load_file(pp)
            for li, k in enumerate(self.local_ids):
                if f"peer_{k}" in saved:
                    self.peer_state[li].copy_(saved[f"peer_{k}"])
            self._peer_state_restored = all(f"peer_{k}" in saved for k in self.local_ids)
        self.log.log("resume", round=self.start_round, path=path, peer_states=bool(getattr(self, "_peer_state_restored", False)))

    def _checkpoint(self, rnd: int) -> Optional[float]:
        """``global_model.save_pretrained(dir)`` every round (serverless_IID_IMDB.py:280-284).  The round loop only takes a
        device-side snapshot; the HF-layout directory is written by a background thread (``--checkpoint-sync`` writes in
        line and adds the cross-rank barrier of round 1).  Gossip topologies also save each rank's per-peer states."""
        cfg = self.cfg
        if not cfg.checkpoint_dir or (rnd + 1) % max(cfg.checkpoint_every, 1) != 0:
            return None
        cuda = self.device.type == "cuda"
        snap = self.global_flat.clone() if self.rank == 0 else None
        peers = self.peer_state[: len(self.local_ids)].clone() if self.peer_state is not None else None
        ev = None
        if cuda:
            ev = torch.cuda.Event()
            ev.record()
        meta = {"round": rnd, "global_accuracies": list(self.global_accuracies), "mode": cfg.mode,
                "ledger_height": self.ledger.height if self.ledger else 0, "masked": list(self.masked)}
        ledger_blob = self.ledger.to_json() if (self.ledger is not None and self.rank == 0) else None
        arena, ids, rank, path, dev = self.runner.arena, list(self.local_ids), self.rank, cfg.checkpoint_dir, self.device

        def job() -> Optional[float]:
            from .models.arena import FlatArena
            if cuda:
                torch.cuda.set_device(dev)
                ev.synchronize()
                if not hasattr(self, "_ck_stream"):
                    self._ck_stream = torch.cuda.Stream(dev)
            os.makedirs(path, exist_ok=True)
            size = None
            with (torch.cuda.stream(self._ck_stream) if cuda else _nullctx()):
                if peers is not None:
                    from safetensors.torch import save_file
                    host = peers.to("cpu")
                    fn = os.path.join(path, f"peer_states_rank{rank}.safetensors")
                    save_file({f"peer_{k}": host[li].contiguous() for li, k in enumerate(ids)}, fn + ".tmp")
                    os.replace(fn + ".tmp", fn)
                if snap is not None:
                    host = snap.to("cpu")
                    state = FlatArena(arena.shapes, torch.float32, "cpu", buffer=host, key_order=arena.key_order).state_dict()
                    size = save_pretrained(state, cfg.model, path, meta)
                    if ledger_blob is not None:
                        with open(os.path.join(path, "ledger.json.tmp"), "w") as fh:
                            fh.write(ledger_blob)
                        os.replace(os.path.join(path, "ledger.json.tmp"), os.path.join(path, "ledger.json"))
            return size

        if cfg.checkpoint_async:
            if self.ckpt_writer is None:
                from .utils.ckpt_writer import AsyncCheckpointWriter
                self.ckpt_writer = AsyncCheckpointWriter()
            self.ckpt_writer.submit(job)
            return self.ckpt_writer.last_size_gib       # size of the newest checkpoint already on disk (may lag a round)
        size = job()
        self.comm.barrier()
        return size

    @staticmethod
    def _raise_on_peer_error(rb: HostReadback, h_err) -> None:
        if h_err is not None and int(rb.get(h_err).view(-1)[0]) != 0:
            raise RuntimeError("a peer-memory barrier timed out during this round (a rank is late or missing): the "
                               "aggregate may be incomplete, so the round is not recorded")

    # ---- one client visit ----------------------------------------------------------------------
    def _train_peer(self, li: int, k: int, rnd: int, start: torch.Tensor, lazy: bool = False) -> Dict:
        """``lazy``: ``loss`` and ``root`` stay device tensors (resolved by the caller with one synchronisation per round)."""
        pin = self.device.type == "cuda"
        tr, te = self._loader_cache.pop((k, rnd), None) or client_loaders(self.corpus, self.cfg.data, k, rnd, pin=pin)
        with self.timer.phase("train"):
            self.runner.load_flat(start)
            self.runner.sync_shadow()
            stats = self.runner.train_epochs(tr, self.cfg.local_epochs, lazy=lazy) if lazy else \
                self.runner.train_epochs(tr, self.cfg.local_epochs)
        corrupted = self.injector.maybe_corrupt(k, self.runner.flat, start, rnd)
        if corrupted and hasattr(self.runner, "mark_dirty"):
            self.runner.mark_dirty()                       # flat was edited in place: the bf16 shadow is stale
        with self.timer.phase("publish"):
            root = self._publish(li, lazy=lazy)
        return {"peer": k, "batches": len(tr), "samples": tr.num_samples, "tokens": tr.num_tokens(),
                "loss": stats["loss"], "root": root, "corrupted": corrupted, "testloader": te}

    def _publish(self, li: int, lazy: bool = False):
        """Copy the trained weights into the peer-visible slot; hash them on the way if the ledger
        is on (fused ``sha256_copy`` kernel on a GPU).  Returns the Merkle root (bytes, or a 32-byte device tensor when
        ``lazy`` on a GPU) or None without a ledger."""
        src, dst = self.runner.flat, self.slots[li]
        if self.ledger is None:
            dst.copy_(src)
            return None
        if src.is_cuda:
            from .ops import cuda_ops
            self.comm.launches += 2
            return cuda_ops.sha256_copy_root(src, dst, self.ledger.chunk_bytes, to_host=not lazy)
        dst.copy_(src)
        from .ledger.merkle import merkle_root
        return merkle_root(src.detach().numpy().tobytes(), self.ledger.chunk_bytes)

    def _score(self, rnd: int) -> List[int]:
        """On-line cosine/PageRank scoring of the published updates -> masked peer ids."""
        if self.cfg.anomaly_detector != "cosine_pagerank":
            return list(self.static_masked)
        with self.timer.phase("score"):
            L = len(self.local_ids)
            on_gpu = self.device.type == "cuda"
            if self.deltas is None:
                self.deltas = self.comm.delta_buffer(self.P, self.D, torch.bfloat16 if on_gpu else torch.float32)
            for li in range(L):
                if on_gpu:
                    from .ops import cuda_ops
                    cuda_ops.delta_bf16(self.slots[li], self.global_flat, self.deltas[li][: self.D])
                    self.comm.launches += 1
                else:
                    torch.sub(self.slots[li], self.global_flat, out=self.deltas[li][: self.D])
            g = self.comm.gram(self.deltas, self.N)
            if self.placement.kind != "block":             # rows come in (rank, slot) order: bring them into peer-id order
                perm = torch.tensor([self.placement.owner(k)[0] * self.P + self.placement.owner(k)[1] for k in range(self.N)],
                                    device=g.device)
                g = g[perm][:, perm]
            if on_gpu and self.N <= 128:
                from .ops import cuda_ops
                ranks_d, mask_d, info_d = cuda_ops.pagerank(g[: self.N, : self.N].contiguous(), mode=1, low_only=True)
                self.comm.launches += 1
                ranks = ranks_d.cpu().numpy()
                mean_rank = float(ranks.mean())          # same guard as anomaly.cosine.score_updates(min_rel_gap=0.1)
                bad = [int(i) for i in np.flatnonzero(mask_d.cpu().numpy()) if abs(ranks[i] - mean_rank) > 0.1 * mean_rank]
                band = (float(info_d[1]), float(info_d[2]))
            else:
                bad, ranks, band = score_updates(g.to(torch.float64).cpu().numpy()[: self.N, : self.N])
        self.log.log("score", round=rnd, masked=bad, ranks=[round(float(x), 6) for x in ranks], band=band)
        return bad

    def _relay_plan(self, w: np.ndarray):
        """Two-hop routes of this round's gossip exchange (path optimiser: Dijkstra on size / bandwidth) turned into an
        executable relay plan; None when routing is off, nothing is routed, or the backend has no links to route over."""
        if not (self.cfg.route_optimizer and self.comm.executes_relays and self.world > 2):
            return None
        key = ("relays", tuple(sorted(self.masked)))
        relays = self._route_cache.get(key)
        if relays is None:
            relays = {}
            for j in range(self.N):
                if j in self.masked:
                    continue
                rp = optimise_route(self.topo, self.nbytes, j, removed=self.masked)
                for i, path in rp.routes.items():
                    if len(path) >= 3:
                        relays[(i, j)] = int(path[-2])          # last hop's sender relays (longer paths: earlier hops direct)
            self._route_cache[key] = relays
        if not relays:
            return None
        plan = RelayPlan(w, self.placement.owner, relays, getattr(self.comm, "relay_slots", 4))
        return plan if plan.any else None

    def _verify_received(self, rb: HostReadback):
        """Receiver-side check of the ledger claim: recompute the Merkle root of the neighbour rank's published slot — the
        bytes this rank's aggregate consumed, read over NVLink — and compare it on the device with the root that rank put into
        the exchanged round record.  Uses the allocation-free entry point (the receive path must not cudaMalloc under the
        other peers' kernels).  Returns the readback handle of the mismatch count."""
        from .utils import records as R
        src_rank = (self.rank + 1) % self.world
        got = torch.empty(32, dtype=torch.uint8, device=self.device)
        if self.device.type == "cuda":
            from .ops import cuda_ops
            ptr = self.comm.peer_ptr(self.slots[0], src_rank)
            if (self.nbytes + self.ledger.chunk_bytes - 1) // self.ledger.chunk_bytes <= 131072:     # FEDB200_STATIC_MERKLE_MAX_LEAVES
                cuda_ops.sha256_root_static(ptr, self.nbytes, self.ledger.chunk_bytes, got)
            else:       # larger than the static workspace (BERT-large at 4 KiB chunks): cached caller-side workspace instead
                from .parallel.peer_mem import _RawCuda
                view = torch.as_tensor(_RawCuda(ptr, self.nbytes), device=self.device).view(torch.float32)
                got = cuda_ops.sha256_copy_root(view, None, self.ledger.chunk_bytes, to_host=False)
            self.comm.launches += 2
        else:
            from .ledger.merkle import merkle_root
            got.copy_(torch.frombuffer(bytearray(merkle_root(self.slots[0].numpy().tobytes(), self.ledger.chunk_bytes)), dtype=torch.uint8))
        want = self.records.table[src_rank, R.R_ROOT:R.R_ROOT + 8].view(torch.uint8)
        return rb.push((got != want).sum().to(torch.int32).reshape(1))

    def _route_of(self, peer: int) -> List[int]:
        """Route recorded in the ledger block of ``peer``'s update.  Gossip topologies: the route that was EXECUTED to the
        peer's slowest mixing neighbour this round (direct, or through the relay GPU the fused mix kernel used).  Full
        topology: the optimiser's path to the slowest destination (the aggregate itself is a direct all-to-all reduce)."""
        if not (self.cfg.route_optimizer and self.ledger is not None):
            return []
        w = getattr(self, "_gossip_w", None)
        if self.peer_state is not None and w is not None:
            hop = self.topo.hop_seconds(self.nbytes)
            dests = [i for i in range(self.N) if i != peer and w[i, peer] != 0.0]
            if not dests:
                return []
            plan = getattr(self, "_gossip_plan", None)
            route_to = lambda i: (plan.executed.get((i, peer), [peer, i]) if plan is not None else [peer, i])
            cost = lambda r: sum(hop[a][b] for a, b in zip(r[:-1], r[1:]))
            return list(max((route_to(i) for i in dests), key=cost))
        key = (peer, tuple(sorted(self.masked)))
        r = self._route_cache.get(key)
        if r is None:
            plan = optimise_route(self.topo, self.nbytes, peer, removed=self.masked)
            r = list(plan.routes[max(plan.routes, key=lambda j: plan.latency[j])]) if plan.routes else []
            self._route_cache[key] = r
        return r

    def _ledger_append(self, rnd: int, records: List[Dict]) -> None:
        """Append this round's blocks (one per peer, every rank appends the same blocks).  ``records`` is the parsed round
        table (:mod:`fedb200.utils.records`) — it came over NVLink inside the round's kernels, no host collective."""
        if self.ledger is None:
            return
        with self.timer.phase("ledger"):
            for r in records:
                self.ledger.append(rnd, r["peer"], BROADCAST, r["root"], self.nbytes, r["batches"], self._route_of(r["peer"]),
                                   timestamp_ns=rnd * 1_000_000 + r["peer"])
            if self.cfg.num_rounds <= 4 or rnd % 5 == 0:
                self.ledger.assert_valid()

    def _aggregate(self, ws: List[float], total: float, overlap: bool) -> None:
        """Global model = weighted mean over every rank's peers.  ``overlap``: issue it on the side stream (SM-light grid) so
        it runs underneath whatever the main stream does next; :meth:`_aggregate_join` orders later work after it."""
        cuda = self.device.type == "cuda"
        side = self._agg_stream if (overlap and cuda and self.world > 1) else None
        if side is not None:
            side.wait_stream(torch.cuda.current_stream())
        with (torch.cuda.stream(side) if side is not None else _nullctx()):
            kw = {"max_ctas": self.cfg.aggregate_ctas} if side is not None else {}
            if self.nvls:
                kw["nvls"] = True
            if self.P == 1 and self.N == self.world and not self.masked:
                # one peer per GPU and nobody masked anywhere: the slot IS the rank's accumulator.  Every rank must take
                # the same branch — the fused reduce reads the SAME symmetric buffer on all peers.
                self.comm.weighted_mean(self.slots[0], 1.0, self.global_flat, **kw)
            else:
                self._local_weighted_sum(ws)
                self.comm.weighted_mean(self.acc, total, self.global_flat, **kw)
        self._agg_pending = side

    def _aggregate_join(self) -> None:
        if getattr(self, "_agg_pending", None) is not None:
            torch.cuda.current_stream().wait_stream(self._agg_pending)
            self._agg_pending = None

    # ---- server mode -----------------------------------------------------------------------------
    def round_server(self, rnd: int) -> Dict:
        from .utils import records as R
        infos = [self._train_peer(li, k, rnd, self.global_flat, lazy=True) for li, k in enumerate(self.local_ids)]
        rb, rec = self.readback, self.records            # one device -> host synchronisation per round (see round_serverless)
        rb.reset()
        rb_bytes0 = rb.bytes_read
        rec.begin(infos)
        for li, i in enumerate(infos):
            rec.set_f32(li, R.R_LOSS, i["loss"])
            rec.set_root(li, i["root"])
        self.masked = self._score(rnd)
        with self.timer.phase("aggregate"):
            ws = [0.0 if i["peer"] in self.masked else float(i["batches"]) for i in infos]   # n_k = len(trainloader)
            self._aggregate(ws, sum(ws), overlap=False)
        with self.timer.phase("evaluate"):
            self.runner.load_flat(self.global_flat)
            self.runner.sync_shadow()
            for li, i in enumerate(infos):
                te = i["testloader"]
                ls, co, n_ds, tot = self.runner.evaluate_lazy(te)
                rec.set_f32(li, R.R_ELOSS, ls)
                rec.set_i32(li, R.R_ECORRECT, co)
                for w, v in ((R.R_ENDS, n_ds), (R.R_ETOTAL, tot), (R.R_EBATCHES, len(te))):
                    rec.set_i32(li, w, v)
        rec.exchange()
        h_tab = rb.push(rec.table)
        ew = self.comm.error_word()
        h_err = rb.push(ew) if ew is not None else None
        rb.sync()
        self._raise_on_peer_error(rb, h_err)             # a timed-out peer barrier fails the round before it is recorded
        self.runner.d2h_bytes = getattr(self.runner, "d2h_bytes", 0) + (rb.bytes_read - rb_bytes0)
        allrec = rec.parse(rb.get(h_tab))
        ev = []
        for r in allrec:
            loss = r["eval_loss_sum"] / max(r["eval_n_ds"], 1)
            acc = r["eval_correct"] / r["eval_total"] if r["eval_total"] else 0.0
            ev.append((float(loss), r["eval_batches"], {"accuracy": float(acc), "loss": float(loss)}))
        self._ledger_append(rnd, allrec)
        loss, metrics = self.strategy.aggregate_evaluate(rnd, ev)
        self.global_accuracies.append(metrics.get("accuracy", float("nan")))
        mine = [r["loss"] for r in allrec if r["peer"] in self.local_ids]
        return {"loss_distributed": loss, **metrics, "train_loss": float(np.mean(mine)) if mine else None,
                "samples": sum(i["samples"] for i in infos), "tokens": sum(i["tokens"] for i in infos)}

    def _local_weighted_sum(self, ws: List[float]) -> None:
        L = len(ws)
        if self.acc.is_cuda:
            from .ops import cuda_ops
            srcs = [self.slots[i] for i in range(L) if ws[i] != 0.0]
            if srcs:
                cuda_ops.mix_n(srcs, [w for w in ws if w != 0.0], self.acc)
                self.comm.launches += 1
            else:
                self.acc.zero_()
        else:
            self.acc.zero_()
            for i in range(L):
                if ws[i] != 0.0:
                    self.acc.add_(self.slots[i], alpha=ws[i])

    # ---- serverless mode ---------------------------------------------------------------------------
    def round_serverless(self, rnd: int) -> Dict:
        from .utils import records as R
        cfg = self.cfg
        infos = []
        if cfg.serverless_semantics == "handoff":
            if self.world != 1:
                raise ValueError("handoff (reference-compat) semantics is a single-process chain")
            cur = self.global_flat.clone()
            for li, k in enumerate(self.local_ids):
                infos.append(self._train_peer(li, k, rnd, cur))
                cur.copy_(self.runner.flat)                  # next client continues from these weights
        else:
            for li, k in enumerate(self.local_ids):
                start = self.global_flat if self.peer_state is None else self.peer_state[li]
                infos.append(self._train_peer(li, k, rnd, start, lazy=True))
        # Device results of this round (training loss, Merkle root, evaluation sums of every peer of every rank) travel in
        # the round-record table: filled with device-side copies, exchanged over NVLink by one small kernel, read back by
        # ONE stream synchronisation after the global evaluation has been issued.
        rb, rec = self.readback, self.records
        rb.reset()
        rb_bytes0 = rb.bytes_read
        rec.begin(infos)
        for li, i in enumerate(infos):
            rec.set_f32(li, R.R_LOSS, i["loss"])
            rec.set_root(li, i["root"])
        online = cfg.anomaly_detector == "cosine_pagerank"
        if not online:
            # the mask is static, so the aggregate depends only on the published slots: start it now on the side stream and
            # let it run underneath the local evaluation (which reads the runner's weights, not the global model)
            self.masked = list(self.static_masked)
            with self.timer.phase("aggregate"):
                ws = [0.0 if i["peer"] in self.masked else 1.0 for i in infos]      # unweighted mean (:269)
                self._aggregate(ws, sum(ws), overlap=self.peer_state is None)
        with self.timer.phase("evaluate_local"):
            for li, i in enumerate(infos):                   # client.evaluate_model() after training (:264)
                if not (len(infos) == 1 and not i["corrupted"] and cfg.serverless_semantics != "handoff"):
                    self.runner.load_flat(self.slots[li])    # with one local peer the runner still holds exactly these
                    self.runner.sync_shadow()                # weights (and AdamW kept the bf16 shadow current)
                ls, co, n_ds, tot = self.runner.evaluate_lazy(i["testloader"])
                rec.set_i32(li, R.R_LCORRECT, co)
                rec.set_i32(li, R.R_LTOTAL, tot)
        if online:
            self.masked = self._score(rnd)
            with self.timer.phase("aggregate"):
                ws = [0.0 if i["peer"] in self.masked else 1.0 for i in infos]
                self._aggregate(ws, sum(ws), overlap=False)
        with self.timer.phase("aggregate"):
            self._aggregate_join()
            if self.peer_state is not None:
                w = apply_mask(mixing_matrix(self.topo.adjacency(), "metropolis"), self.masked)
                rows = w[self.local_ids] if self.local_ids else w[:0]
                if rows.shape[0] < self.P:
                    rows = np.concatenate([rows, np.zeros((self.P - rows.shape[0], self.N))], 0)
                plan = self._relay_plan(w)
                self.comm.mix_rows(self.slots, rows, self.peer_state, self.N, plan=plan)
                self._gossip_plan, self._gossip_w = plan, w
        rec.exchange()
        h_ver = self._verify_received(rb) if (cfg.verify_updates and self.ledger is not None) else None
        with self.timer.phase("evaluate_global"):
            self.runner.load_flat(self.global_flat)
            self.runner.sync_shadow()
            te = global_test_loader(self.corpus, cfg.data, rnd, pin=self.device.type == "cuda")
            g_ls, g_co, g_nds, g_tot = self.runner.evaluate_lazy(te)
            h_gl, h_gc = rb.push(g_ls.to(torch.float32)), rb.push(g_co.to(torch.int64))
        # next round's loaders are built now, while the GPU is still busy with the global evaluation
        self._loader_cache.clear()
        if rnd + 1 < cfg.num_rounds:
            for k in self.local_ids:
                self._loader_cache[(k, rnd + 1)] = client_loaders(self.corpus, cfg.data, k, rnd + 1, pin=self.device.type == "cuda")
        h_tab = rb.push(rec.table)
        ew = self.comm.error_word()
        h_err = rb.push(ew) if ew is not None else None
        rb.sync()                                            # the round's only device -> host synchronisation
        self._raise_on_peer_error(rb, h_err)                 # a timed-out peer barrier fails the round before it is recorded
        if h_ver is not None and int(rb.get(h_ver).view(-1)[0]) != 0:
            raise RuntimeError("a received update does not hash to the Merkle root its sender recorded (round not recorded)")
        self.runner.d2h_bytes = getattr(self.runner, "d2h_bytes", 0) + (rb.bytes_read - rb_bytes0)
        allrec = rec.parse(rb.get(h_tab))
        gl = float(rb.get(h_gl)) / g_nds
        ga = int(rb.get(h_gc)) / g_tot if g_tot else 0.0
        self._ledger_append(rnd, allrec)
        locs = [(r["local_correct"] / r["local_total"] if r["local_total"] else 0.0) for r in allrec]
        mine = [r["loss"] for r in allrec if r["peer"] in self.local_ids]
        self.global_accuracies.append(float(ga))
        return {"accuracy": float(ga), "loss": float(gl), "local_accuracy": locs,
                "train_loss": float(np.mean(mine)) if mine else None,
                "samples": sum(i["samples"] for i in infos), "tokens": sum(i["tokens"] for i in infos)}

    # ---- asynchronous serverless (virtual-time event engine) -------------------------------------------
    def run_async_scheduled(self) -> Dict:
        """Multi-GPU asynchronous gossip: every rank derives the same virtual-time schedule
        (:mod:`fedb200.p2p.schedule`) and executes its own peers' events; cross-GPU ordering is enforced by
        device-side progress flags in peer memory (``pub`` / ``done`` epochs), never by host barriers."""
        from .p2p import build_async_schedule, model_durations
        from .parallel.placement import owner_of
        cfg, N, D, comm = self.cfg, self.N, self.D, self.comm
        on_gpu = self.device.type == "cuda"
        if self.world > 1 and comm.name != "peer":
            raise NotImplementedError("multi-rank asynchronous gossip needs the peer-memory backend (device flags)")
        masked = list(self.static_masked)
        dur = model_durations(N, cfg.async_epoch_s, cfg.async_spread, cfg.seed)
        sched = build_async_schedule(self.topo, self.nbytes, cfg.num_rounds, dur, cfg.staleness_decay, masked=masked,
                                     route_optimizer=cfg.route_optimizer, fanout=cfg.async_fanout,
                                     min_latency=cfg.async_min_latency_s)
        R, P = sched.slots_per_peer, self.P
        outbox = comm.alloc("async_outbox", P * R * D).view(P, R, D)
        state = comm.alloc("async_state", P * D).view(P, D)
        flags = comm.alloc("async_flags", 2 * N + 2, torch.int32)        # pub[0, N) | done[N, 2N)
        for li in range(len(self.local_ids)):
            state[li].copy_(self.global_flat)
        if on_gpu:
            torch.cuda.synchronize()
        comm.barrier()
        mine = [ev for ev in sched.events if owner_of(ev.peer, N, self.world)[0] == self.rank]
        root_table = torch.zeros(max(len(mine), 1), 32, dtype=torch.uint8, device=self.device) if self.ledger is not None else None
        if self.ledger is not None and on_gpu:
            from .ops import cuda_ops
            nch = (self.nbytes + self.ledger.chunk_bytes - 1) // self.ledger.chunk_bytes
            leaves_ws = torch.empty(nch, 32, dtype=torch.uint8, device=self.device)
            scratch_ws = torch.empty((nch + 1) // 2 + 1, 32, dtype=torch.uint8, device=self.device)
        samples = tokens = 0
        t0 = time.perf_counter()
        for n_mine, ev in enumerate(mine):
            li = owner_of(ev.peer, N, self.world)[1]
            tr, _ = client_loaders(self.corpus, cfg.data, ev.peer, ev.epoch - 1, pin=on_gpu)
            with self.timer.phase("train"):
                self.runner.load_flat(state[li])
                self.runner.sync_shadow()
                self.runner.train_epochs(tr, cfg.local_epochs)
                self.injector.maybe_corrupt(ev.peer, self.runner.flat, state[li], ev.epoch - 1)
            samples, tokens = samples + tr.num_samples, tokens + tr.num_tokens()
            with self.timer.phase("aggregate"):
                srcs, ws = [self.runner.flat], [ev.self_weight]
                for s_peer, s_epoch, w in zip(ev.sources, ev.source_epochs, ev.weights):
                    rs, ls = owner_of(s_peer, N, self.world)
                    if rs != self.rank:
                        comm.flag_wait(flags, rs, s_peer, s_epoch)            # that version has been published
                    slot = sched.slot_of(s_peer, s_epoch)
                    srcs.append(comm.peer_ptr(outbox[ls, slot], rs) if on_gpu else outbox[ls, slot])
                    ws.append(w)
                if on_gpu:
                    from .ops import cuda_ops
                    cuda_ops.mix_n(srcs, ws, state[li])
                    comm.launches += 1
                else:
                    acc = torch.zeros_like(state[li])
                    for b, w in zip(srcs, ws):
                        acc.add_(b, alpha=float(w))
                    state[li].copy_(acc)
                comm.flag_set(flags, N + ev.peer, ev.epoch)                   # done: sources may be overwritten
            with self.timer.phase("publish"):
                for c_peer, c_epoch in ev.overwrite_waits:                     # consumers of the version this slot held
                    rc = owner_of(c_peer, N, self.world)[0]
                    if rc != self.rank:
                        comm.flag_wait(flags, rc, N + c_peer, c_epoch)
                dst = outbox[li, ev.slot]
                if self.ledger is not None and on_gpu:
                    cuda_ops.sha256_leaves(state[li], dst, self.ledger.chunk_bytes, leaves_ws)
                    cuda_ops.merkle_root(leaves_ws, scratch_ws, root_table[n_mine])
                    comm.launches += 2
                else:
                    dst.copy_(state[li])
                    if self.ledger is not None:
                        from .ledger.merkle import merkle_root
                        root_table[n_mine] = torch.frombuffer(bytearray(merkle_root(dst.numpy().tobytes(), self.ledger.chunk_bytes)),
                                                              dtype=torch.uint8)
                comm.flag_set(flags, ev.peer, ev.epoch)                        # pub
        if on_gpu:
            torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        comm.barrier()
        comm.check_errors()
        # ---- consensus model, evaluation, ledger ----------------------------------------------------------------
        keep_w = [0.0 if k in masked else 1.0 for k in self.local_ids]
        L = len(self.local_ids)
        if on_gpu:
            from .ops import cuda_ops
            live = [i for i in range(L) if keep_w[i] != 0.0]
            if live:
                cuda_ops.mix_n([state[i] for i in live], [1.0] * len(live), self.acc)
            else:
                self.acc.zero_()
        else:
            self.acc.zero_()
            for i in range(L):
                if keep_w[i]:
                    self.acc.add_(state[i])
        comm.weighted_mean(self.acc, float(sum(keep_w)), self.global_flat)
        self.runner.load_flat(self.global_flat)
        self.runner.sync_shadow()
        gl, ga = self.runner.evaluate(global_test_loader(self.corpus, cfg.data, cfg.num_rounds, pin=on_gpu))
        self.global_accuracies.append(float(ga))
        if self.ledger is not None:
            rows = root_table.cpu().numpy()
            local = [(ev.index, ev.peer, ev.epoch, bytes(rows[i]).hex(), ev.dests[:16]) for i, ev in enumerate(mine)]
            for idx, peer, epoch, root_hex, dests in sorted(sum(comm.all_gather_object(local), [])):
                self.ledger.append(epoch, peer, BROADCAST, bytes.fromhex(root_hex), self.nbytes, 0, dests,
                                   timestamp_ns=idx)
            self.ledger.assert_valid()
        tot = comm.reduce_scalars([samples, tokens], "sum")
        mx = comm.reduce_scalars([wall], "max")[0]
        res = {"accuracy": float(ga), "loss": float(gl), "mixes": len(sched.events), "makespan_model_s": sched.makespan,
               "async_comm_time": sched.async_comm_time, "sync_equiv_comm_time": sched.sync_equiv_comm_time,
               "outbox_slots": R, "masked": masked, "wall_s": mx, "samples_per_s": tot[0] / mx if mx else None,
               "info_passing": info_passing_time(self.topo, self.nbytes, 0, removed=masked)}
        self.async_schedule = sched
        self.async_state = state
        res["phases_ms"] = self.timer.device_ms()        # this rank's device time per phase (aggregate includes flag waits)
        self.log.log("async_done", **res)
        return res

    def run_async(self) -> Dict:
        if self.cfg.async_scheduled or self.cfg.async_mode == "scheduled":
            return self.run_async_scheduled()
        if self.cfg.async_mode == "observed" or (self.world != 1 and self.cfg.async_mode != "engine"):
            from .p2p.observed import run_async_observed
            return run_async_observed(self)
        cfg = self.cfg
        state = self.comm.alloc("async_state", self.N * self.D).view(self.N, self.D)
        inbox = {}
        for k in range(self.N):
            state[k].copy_(self.global_flat)
        epochs = np.zeros(self.N, dtype=np.int64)
        modelled = None
        if cfg.async_time == "modelled":                   # deterministic virtual clock (same model as the scheduler)
            from .p2p import model_durations
            modelled = model_durations(self.N, cfg.async_epoch_s, cfg.async_spread, cfg.seed)

        def train_fn(k: int) -> float:
            rnd = int(epochs[k])
            tr, _ = client_loaders(self.corpus, cfg.data, k, rnd, pin=self.device.type == "cuda")
            t0 = time.perf_counter()
            with self.timer.phase("train"):
                self.runner.load_flat(state[k])
                self.runner.sync_shadow()
                self.runner.train_epochs(tr, cfg.local_epochs)
                self.injector.maybe_corrupt(k, self.runner.flat, state[k], rnd)
                state[k].copy_(self.runner.flat)
            epochs[k] += 1
            if modelled is not None:
                return float(modelled[k])
            if self.device.type == "cuda":
                torch.cuda.synchronize()
            return time.perf_counter() - t0

        # Data plane: one snapshot per published VERSION (src, epoch), shared by all its destinations and reference
        # counted; an inbox entry (dst, src) names the version that has LANDED there (bound by land_fn), so a source that
        # republishes before its previous version arrived cannot change what a logged mix reads.
        versions: Dict = {}                               # (src, epoch) -> [tensor, root, references]
        self.async_audit: List = []                       # filled when self._audit_async is set (tests)

        def release(key) -> None:
            v = versions.get(key)
            if v is not None:
                v[2] -= 1
                if v[2] <= 0:
                    del versions[key]

        def publish_fn(src: int, dst: int) -> None:
            key = (src, int(epochs[src]))
            with self.timer.phase("publish"):
                v = versions.get(key)
                if v is None:
                    buf = torch.empty_like(state[src])
                    if self.ledger is not None and buf.is_cuda:
                        from .ops import cuda_ops
                        root = cuda_ops.sha256_copy_root(state[src], buf, self.ledger.chunk_bytes)
                        self.comm.launches += 2
                    else:
                        buf.copy_(state[src])
                        root = None
                        if self.ledger is not None:
                            from .ledger.merkle import merkle_root
                            root = merkle_root(buf.numpy().tobytes(), self.ledger.chunk_bytes)
                    v = versions[key] = [buf, root, 0]
                    if getattr(self, "_audit_async", False):
                        self.async_audit.append(("pub", src, key[1], float(buf.double().sum())))
                v[2] += 1
                if self.ledger is not None:
                    self.ledger.append(key[1], src, dst, v[1], self.nbytes, 0, [src, dst])

        def land_fn(dst: int, src: int, epoch: int, accepted: bool) -> None:
            if not accepted:
                release((src, epoch))
                return
            old = inbox.get((dst, src))
            if old is not None:
                release((src, old))
            inbox[(dst, src)] = epoch

        def mix_fn(k: int, srcs: List[int], ws: List[float], self_w: float) -> None:
            with self.timer.phase("aggregate"):
                if srcs:
                    used = [(s, inbox[(k, s)]) for s in srcs]
                    bufs = [state[k]] + [versions[u][0] for u in used]
                    wts = [self_w] + list(ws)
                    if getattr(self, "_audit_async", False):
                        self.async_audit.append(("mix", k, int(epochs[k]), [(s, e, float(versions[(s, e)][0].double().sum()))
                                                                             for s, e in used]))
                    if state.is_cuda:
                        from .ops import cuda_ops
                        cuda_ops.mix_n(bufs, wts, self.acc)
                        self.comm.launches += 1
                        state[k].copy_(self.acc)
                    else:
                        self.acc.zero_()
                        for b, w in zip(bufs, wts):
                            self.acc.add_(b, alpha=float(w))
                        state[k].copy_(self.acc)
                for (d, s) in [key for key in inbox if key[0] == k]:        # the engine clears k's inbox after every mix
                    release((s, inbox.pop((d, s))))

        masked = list(self.static_masked)
        eng = AsyncGossipEngine(self.topo, self.nbytes, train_fn, mix_fn, publish_fn, cfg.staleness_decay,
                                masked=masked, route_optimizer=cfg.route_optimizer, min_latency=cfg.async_min_latency_s,
                                fanout=cfg.async_fanout, land_fn=land_fn)
        res = eng.run(cfg.num_rounds)
        keep = [k for k in range(self.N) if k not in masked]
        self.global_flat.zero_()
        for k in keep:
            self.global_flat.add_(state[k], alpha=1.0 / len(keep))
        self.runner.load_flat(self.global_flat)
        self.runner.sync_shadow()
        gl, ga = self.runner.evaluate(global_test_loader(self.corpus, cfg.data, cfg.num_rounds))
        self.global_accuracies.append(float(ga))
        res.update(accuracy=float(ga), loss=float(gl), masked=masked,
                   info_passing=info_passing_time(self.topo, self.nbytes, 0, removed=masked))
        self.async_engine = eng
        self.log.log("async_done", **res)
        return res

    # ---- driver ----------------------------------------------------------------------------------------
    def run(self) -> Dict:
        cfg = self.cfg
        if cfg.mode == "serverless" and cfg.asynchronous:
            out = self.run_async()
            return self._finish(out)
        last = {}
        for rnd in range(self.start_round, cfg.num_rounds):
            t0 = time.perf_counter()
            self.timer.reset()
            last = self.round_server(rnd) if cfg.mode == "server" else self.round_serverless(rnd)
            size = self._checkpoint(rnd)
            dev = self.timer.device_ms()
            wall = time.perf_counter() - t0
            tot = self.comm.reduce_scalars([last.get("samples", 0), last.get("tokens", 0)], "sum")
            mx = self.comm.reduce_scalars([wall], "max")[0]
            # bytes this rank moves over NVLink in the fused two-shot FedAvg (reduce-scatter reads + all-gather pushes) and
            # the fraction of the link roofline (770 GB/s per direction, measured) the aggregate phase reached; the phase
            # also contains the wait for the slowest rank, so the fraction is a lower bound
            W = self.world
            comm_bytes = int(2 * (W - 1) / W * self.nbytes) if W > 1 else 0
            agg_ms = dev.get("aggregate", 0.0)
            # (None when the aggregate ran on the side stream: the main-stream phase timer then only sees its join)
            nvlink_frac = (comm_bytes / 2 / 770e9 * 1e3) / agg_ms if comm_bytes and agg_ms > 0.2 else None
            self.log.log("round", round=rnd, wall_s=mx, samples_per_s=tot[0] / mx if mx else None,
                         tokens_per_s=tot[1] / mx if mx else None, phases_ms=dev, model_size_gib=size,
                         comm_bytes_per_rank=comm_bytes, aggregate_nvlink_frac=nvlink_frac,
                         masked=self.masked, ledger_height=self.ledger.height if self.ledger else 0,
                         **{k: v for k, v in last.items() if k not in ("samples", "tokens")})
        return self._finish(last)

    def _finish(self, last: Dict) -> Dict:
        if self.ckpt_writer is not None:
            last = dict(last, model_size_gib=self.ckpt_writer.flush())       # newest checkpoint is on disk from here on
            self.comm.barrier()
        rep = self.overhead.report()
        comm_model = info_passing_time(self.topo, self.nbytes, 0, removed=self.masked)
        out = {"global_accuracies": self.global_accuracies, **rep, "info_passing_time": comm_model,
               "ledger_height": self.ledger.height if self.ledger else 0, "last": last}
        self.log.log("done", **out)
        return out

    def close(self) -> None:
        if self.ckpt_writer is not None:
            self.ckpt_writer.close()
            self.ckpt_writer = None
        self.log.close()
        if self._own_comm:
            self.comm.close()


def run_experiment(cfg: FedConfig, echo: bool = True) -> Dict:
    rt = FedRuntime(cfg, echo=echo)
    try:
        return rt.run()
    finally:
        rt.close()


def run_client_sweep(cfg: FedConfig, counts=(5, 10, 20), echo: bool = True) -> Dict[int, Dict]:
    """Client-count sweep (C16): ``for NUM_CLIENTS in [5, 10, 20]`` re-running the experiment
    (src/Serverlesscase/serverless_cancer_biobert_allclients.py:40-41)."""
    import copy
    out = {}
    for n in counts:
        c = copy.deepcopy(cfg)
        c.num_clients = n
        out[n] = run_experiment(c, echo=echo)
    return out
