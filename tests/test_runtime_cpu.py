"""End-to-end CPU plumbing: the four entrypoints, resume, sweep, async, handoff semantics, 2-process gloo."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from fedb200.config import FedConfig, ModelConfig
from fedb200.runtime import FedRuntime, run_client_sweep, run_experiment

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cfg(mode="serverless", **kw):
    cfg = FedConfig(mode=mode, num_clients=3, num_rounds=2, device="cpu", backend="torch")
    cfg.model = ModelConfig.preset("bert-tiny", num_labels=2)
    cfg.data.population, cfg.data.train_per_client, cfg.data.test_per_client = 600, 32, 32
    cfg.data.max_seq_len, cfg.lr = 48, 2e-3
    for k, v in kw.items():
        setattr(cfg, k, v)
    return cfg


def test_serverless_rounds_ledger_and_checkpoint(tmp_path):
    cfg = _cfg(ledger=True, checkpoint_dir=str(tmp_path / "ck"), metrics_path=str(tmp_path / "m.jsonl"))
    out = run_experiment(cfg, echo=False)
    assert len(out["global_accuracies"]) == 2 and out["ledger_height"] == 6
    assert {"cpu_overhead_pct", "memory_usage_gb", "latency_min", "info_passing_time"} <= set(out)
    assert os.path.exists(tmp_path / "ck" / "config.json") and os.path.exists(tmp_path / "ck" / "ledger.json")
    recs = [json.loads(l) for l in open(tmp_path / "m.jsonl")]
    rounds = [r for r in recs if r["event"] == "round"]
    assert len(rounds) == 2 and rounds[0]["ledger_height"] == 3 and "train" in rounds[0]["phases_ms"]
    # resume continues at round 2 and keeps the validated ledger
    cfg2 = _cfg(ledger=True, checkpoint_dir=str(tmp_path / "ck"), resume=True, num_rounds=3)
    rt = FedRuntime(cfg2, echo=False)
    assert rt.start_round == 2 and rt.ledger.height == 6
    out2 = rt.run()
    rt.close()
    assert len(out2["global_accuracies"]) == 3 and out2["ledger_height"] == 9


def test_server_mode_learns_and_masks_injected_anomaly():
    cfg = _cfg("server", num_clients=5, num_rounds=3, anomaly_detector="cosine_pagerank", inject_anomalies=1,
               inject_kind="signflip")
    cfg.data.train_per_client = 64
    rt = FedRuntime(cfg, echo=False)
    target = rt.injector.targets()
    out = rt.run()
    rt.close()
    assert rt.masked == target
    assert len(out["global_accuracies"]) == 3 and all(0.0 <= a <= 1.0 for a in out["global_accuracies"])


def test_handoff_semantics_matches_reference_chain():
    """With DEVICE=cpu the reference's snapshots alias the live weights, so its 'mean' equals the final
    sequentially-trained weights; our handoff mode averages real copies along the chain."""
    cfg = _cfg(serverless_semantics="handoff", num_rounds=1)
    rt = FedRuntime(cfg, echo=False)
    g0 = rt.global_flat.clone()
    rt.round_serverless(0)
    mean = rt.slots[:3].mean(0)
    torch.testing.assert_close(rt.global_flat, mean)
    chain = rt.slots[:3].clone()
    rt.close()
    par = FedRuntime(_cfg(serverless_semantics="parallel", num_rounds=1), echo=False)
    par.round_serverless(0)
    torch.testing.assert_close(par.slots[0], chain[0])          # the first client starts from the global model in both
    assert (par.slots[1] - chain[1]).abs().max().item() > 1e-6   # later clients continued from their predecessor
    assert (par.slots[2] - chain[2]).abs().max().item() > 1e-6
    par.close()


def test_async_serverless_and_sparse_topology():
    cfg = _cfg(asynchronous=True, topology="reference10", num_clients=4, num_rounds=2, ledger=True)
    out = run_experiment(cfg, echo=False)
    assert out["last"]["mixes"] == 8 and out["last"]["async_comm_time"] <= out["last"]["sync_equiv_comm_time"]
    assert out["ledger_height"] > 0
    cfg = _cfg(topology="ring", num_clients=4, num_rounds=2)
    out = run_experiment(cfg, echo=False)
    assert len(out["global_accuracies"]) == 2


def test_static_graph_detector_masks_reference_nodes():
    cfg = _cfg(num_clients=10, num_rounds=1, topology="reference10", anomaly_detector="pagerank")
    cfg.data.train_per_client = cfg.data.test_per_client = 8
    rt = FedRuntime(cfg, echo=False)
    rt.run()
    assert rt.masked == [0, 4, 7, 9]
    rt.close()


def test_client_sweep():
    cfg = _cfg(num_rounds=1)
    res = run_client_sweep(cfg, (2, 3), echo=False)
    assert set(res) == {2, 3} and all(len(r["global_accuracies"]) == 1 for r in res.values())


@pytest.mark.parametrize("script,extra", [("server_IID.py", []), ("server_NonIID.py", ["--split", "dirichlet"]),
                                          ("serverless_IID.py", ["--ledger"]), ("serverless_NonIID.py", [])])
def test_entrypoints_run(script, extra):
    cmd = [sys.executable, os.path.join(ROOT, script), "--model", "bert-tiny", "--clients", "2", "--rounds", "1",
           "--train-per-client", "16", "--test-per-client", "16", "--max-seq-len", "32", "--device", "cpu", "--quiet"] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert len(out["global_accuracies"]) == 1


def test_two_process_gloo_server_round_matches_single_process():
    """world_size 2 over gloo: clients spread over ranks, aggregation through the torch.distributed backend."""
    code = r"""
import os, sys, json, torch
sys.path.insert(0, %r)
from fedb200.config import FedConfig, ModelConfig
from fedb200.runtime import FedRuntime
cfg = FedConfig(mode="server", num_clients=4, num_rounds=1, device="cpu", backend="torch", comm="nccl")
cfg.model = ModelConfig.preset("bert-tiny", num_labels=2, hidden_dropout=0.0, attention_dropout=0.0)
cfg.data.population, cfg.data.train_per_client, cfg.data.test_per_client, cfg.data.max_seq_len = 400, 16, 16, 32
rt = FedRuntime(cfg, echo=False)
rt.run()
if rt.rank == 0:
    print("CHK", float(rt.global_flat.double().sum()), float(rt.global_flat.double().abs().sum()))
rt.close()
""" % ROOT
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
    r2 = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                         "--master-addr", "127.0.0.1", "--master-port", "29533", "-c", code] if False else
                        [sys.executable, "-c", "pass"], capture_output=True, text=True)
    path = os.path.join(ROOT, "gpurun_out", "_gloo_probe.py")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as fh:
        fh.write(code)
    r2 = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                         "--master-addr", "127.0.0.1", "--master-port", "29533", path], capture_output=True, text=True,
                        timeout=900, env=env, cwd=ROOT)
    assert r2.returncode == 0, r2.stderr[-3000:]
    r1 = subprocess.run([sys.executable, path], capture_output=True, text=True, timeout=900,
                        env={k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE")}, cwd=ROOT)
    assert r1.returncode == 0, r1.stderr[-3000:]
    a = [float(x) for x in [l for l in r2.stdout.splitlines() if l.startswith("CHK")][0].split()[1:]]
    b = [float(x) for x in [l for l in r1.stdout.splitlines() if l.startswith("CHK")][0].split()[1:]]
    assert a == pytest.approx(b, rel=1e-5)


def test_scheduled_async_executor_matches_schedule_and_ledger():
    """Schedule-driven async executor (the multi-GPU path) on the loopback backend: every scheduled mix and publish
    happens once, slots are reused exactly, the ledger records every publish in global order."""
    from fedb200.p2p import build_async_schedule, model_durations
    cfg = _cfg(asynchronous=True, async_scheduled=True, num_clients=5, num_rounds=3, ledger=True)
    rt = FedRuntime(cfg, echo=False)
    out = rt.run()
    sched = rt.async_schedule
    assert out["last"]["mixes"] == 15 == len(sched.events)
    assert rt.ledger.height == 15 and rt.ledger.validate() == -1
    assert [b.sender for b in rt.ledger.blocks] == [ev.peer for ev in sched.events]
    assert out["last"]["async_comm_time"] <= out["last"]["sync_equiv_comm_time"]
    # schedule invariants: weights sum to one, a slot is only reused after its last consumer
    for ev in sched.events:
        assert abs(sum(ev.weights) + ev.self_weight - 1.0) < 1e-12
        for c_peer, c_epoch in ev.overwrite_waits:
            c_idx = next(e.index for e in sched.events if e.peer == c_peer and e.epoch == c_epoch)
            assert c_idx < ev.index
    again = build_async_schedule(rt.topo, rt.nbytes, 3, model_durations(5, cfg.async_epoch_s, cfg.async_spread, cfg.seed),
                                 cfg.staleness_decay)
    assert [(e.peer, e.epoch, e.sources, e.slot) for e in again.events] == [(e.peer, e.epoch, e.sources, e.slot) for e in sched.events]
    rt.close()


def test_host_readback_roundtrip():
    """HostReadback: queued tensors of mixed dtype/shape come back bit-exact after one sync; reset() recycles the buffer."""
    import torch
    from fedb200.utils.readback import HostReadback
    rb = HostReadback("cpu", capacity=256)
    a, b, c = torch.tensor(3.25), torch.arange(32, dtype=torch.uint8), torch.tensor(-7, dtype=torch.int64)
    ha, hb, hc = rb.push(a), rb.push(b), rb.push(c)
    rb.sync()
    assert float(rb.get(ha)) == 3.25 and int(rb.get(hc)) == -7 and torch.equal(rb.get(hb), b)
    assert rb.bytes_read == 4 + 32 + 8
    rb.reset()
    assert rb.off == 0
    import pytest
    with pytest.raises(RuntimeError):
        for _ in range(20):
            rb.push(b)


def test_resume_on_ring_topology_restores_per_peer_states(tmp_path):
    """Gossip topologies keep one state per peer; a resumed run must continue from those states (and, without them, from the
    checkpointed consensus) — never from the initial weights under a continued ledger."""
    cfg = _cfg(topology="ring", ledger=True, checkpoint_dir=str(tmp_path / "ck"), num_rounds=2, num_clients=5)
    rt = FedRuntime(cfg, echo=False)
    init = rt.global_flat.clone()
    rt.run()
    states = rt.peer_state.clone()
    consensus = rt.global_flat.clone()
    rt.close()
    assert (tmp_path / "ck" / "peer_states_rank0.safetensors").exists()
    assert float((states[0] - init).abs().max()) > 0 and float((states[0] - states[1]).abs().max()) > 0
    rt2 = FedRuntime(_cfg(topology="ring", ledger=True, checkpoint_dir=str(tmp_path / "ck"), resume=True, num_rounds=3,
                          num_clients=5), echo=False)
    assert rt2.start_round == 2
    torch.testing.assert_close(rt2.peer_state, states)
    torch.testing.assert_close(rt2.global_flat, consensus)
    out = rt2.run()
    rt2.close()
    assert len(out["global_accuracies"]) == 3 and out["ledger_height"] == 15
    # per-peer file missing (older checkpoint): peers restart from the checkpointed consensus, not from the init weights
    os.remove(tmp_path / "ck" / "peer_states_rank0.safetensors")
    rt3 = FedRuntime(_cfg(topology="ring", checkpoint_dir=str(tmp_path / "ck"), resume=True, num_rounds=4, num_clients=5),
                     echo=False)
    torch.testing.assert_close(rt3.peer_state[1], rt3.global_flat)
    assert float((rt3.global_flat - init).abs().max()) > 0
    rt3.close()


def test_async_single_process_mixes_the_version_the_log_names_when_latency_exceeds_an_epoch():
    """Visibility lag > one epoch: a source republishes before its previous version has landed.  Every mix must read exactly
    the (source, epoch) versions its log record names (replayable event log)."""
    cfg = _cfg(asynchronous=True, num_clients=3, num_rounds=6, async_time="modelled", async_epoch_s=0.05,
               async_min_latency_s=0.12)
    rt = FedRuntime(cfg, echo=False)
    rt._audit_async = True
    rt.run()
    eng, audit = rt.async_engine, rt.async_audit
    rt.close()
    pubs = {(src, ep): chk for kind, src, ep, chk in [a for a in audit if a[0] == "pub"]}
    mixes = [a for a in audit if a[0] == "mix"]
    logged = [r for r in eng.log if r.sources]
    assert len(mixes) == len(logged) and len(mixes) > 0
    stale = 0
    for (_, peer, epoch, used), rec in zip(mixes, logged):
        assert peer == rec.peer and epoch == rec.epoch
        assert [(s, e) for s, e, _ in used] == list(zip(rec.sources, rec.source_epochs))
        for s, e, chk in used:
            assert chk == pubs[(s, e)], "mixed bytes are not the logged version"
            stale += int(e < epoch - 1)
    assert stale > 0, "the lag never exceeded an epoch; the test does not exercise the overwrite hazard"


def test_async_checkpoint_writer_latest_wins_and_all_masked_keeps_model(tmp_path):
    from fedb200.parallel.comm import LoopbackComm
    cfg = _cfg(checkpoint_dir=str(tmp_path / "ck"), num_rounds=3)
    assert cfg.checkpoint_async
    rt = FedRuntime(cfg, echo=False)
    out = rt.run()
    w = rt.ckpt_writer
    assert w.written >= 1 and w.written + w.dropped == 3 and out["last"]["model_size_gib"] > 0
    from fedb200.models.hf_io import load_state
    sd = load_state(str(tmp_path / "ck"))                  # flushed by run(): the newest round is on disk
    from fedb200.models.arena import FlatArena
    want = FlatArena(rt.runner.arena.shapes, buffer=rt.global_flat.clone(), key_order=rt.runner.arena.key_order).state_dict()
    for k, v in want.items():
        torch.testing.assert_close(sd[k], v)
    g = rt.global_flat.clone()
    rt.comm.weighted_mean(rt.acc, 0.0, rt.global_flat)     # every contributor masked: the global model is kept
    torch.testing.assert_close(rt.global_flat, g)
    rt.close()


def test_async_observed_mode_replays_from_logged_versions():
    """Observed-version asynchronous gossip (CPU twin of the K3 device path): every mix must equal the host replay from the
    versions it logged, the fan-in is bounded, later epochs mix the sources' latest published versions."""
    from fedb200.p2p.observed import replay_mix
    cfg = _cfg(asynchronous=True, async_mode="observed", num_clients=5, num_rounds=3, ledger=True, async_fanin=2,
               topology="reference10" if False else "full")
    rt = FedRuntime(cfg, echo=False)
    rt._audit_async = True
    out = rt.run()
    recs, aud, table = rt.async_records, rt.async_audit, rt.async_table
    rt.close()
    assert out["last"]["mixes"] == 15 and out["last"]["torn_chunks"] == 0 and out["ledger_height"] == 15
    assert all(len(r["sources"]) == 2 for r in recs)
    assert any(v > 0 for r in recs for v in r["versions"])
    for r in recs:
        want = replay_mix(aud["trained"][(r["peer"], r["epoch"])], r["versions"], aud["published"], r["sources"],
                          r["epoch"], table)
        torch.testing.assert_close(aud["published"][(r["peer"], r["epoch"])], want, rtol=1e-6, atol=1e-7)
        assert all(0 <= v <= r["epoch"] for v in r["versions"])


def test_receiver_verifies_published_update_against_recorded_root():
    """--verify-updates: the consumer recomputes the Merkle root of the slot it aggregates and compares it with the root in the
    exchanged round record; a slot that was altered after hashing fails the round."""
    cfg = _cfg(ledger=True, verify_updates=True, num_clients=1, num_rounds=1)
    rt = FedRuntime(cfg, echo=False)
    rt.round_serverless(0)                                   # honest: passes
    orig = rt._publish

    def tampering_publish(li, lazy=False):
        root = orig(li, lazy=lazy)
        rt.slots[li][7] += 1.0                               # bytes served differ from the bytes hashed
        return root
    rt._publish = tampering_publish
    with pytest.raises(RuntimeError, match="does not hash to the Merkle root"):
        rt.round_serverless(1)
    rt.close()


def test_deterministic_flag_parses_and_cpu_rounds_reproduce_their_ledger_roots():
    """--deterministic reaches the config, and two runs from the same seed record the same Merkle roots (on the CPU / PyTorch
    path that is torch's deterministic mode; the sm_100a path has tests/test_gpu_fed.py::test_deterministic_rounds_...)."""
    from fedb200.config import build_parser, config_from_args
    cfg = config_from_args(build_parser().parse_args(["--deterministic", "--clients", "3"]))
    assert cfg.deterministic is True and cfg.num_clients == 3

    def roots():
        rt = FedRuntime(_cfg(ledger=True, deterministic=True), echo=False)
        try:
            rt.run()
            return [b.update_root for b in rt.ledger.blocks], rt.runner.flat.clone()
        finally:
            rt.close()
            torch.use_deterministic_algorithms(False)
    (r1, w1), (r2, w2) = roots(), roots()
    assert len(r1) == 6 and r1 == r2 and torch.equal(w1, w2)
