"""FedAvg, gossip mixing, anomaly detectors vs the notebook's recorded outputs, path optimiser, async engine, ledger."""
import hashlib

import numpy as np
import pytest
import torch

from fedb200.anomaly import (AnomalyInjector, community_anomalies, cosine_gram, dbscan_anomalies, detect,
                             inject_anomaly, modz_anomalies, pagerank, pagerank_anomalies, score_updates)
from fedb200.ledger import BlockHeader, Ledger, LedgerError, merkle_root
from fedb200.ledger.merkle import leaf_digests, reduce_level
from fedb200.p2p import (AsyncGossipEngine, Topology, apply_mask, info_passing_time, mix_flat, mixing_matrix,
                         optimise_route, staleness_weight)
from fedb200.p2p.path_opt import shortest_paths
from fedb200.server import FedAvg, ServerConfig, aggregate_weighted, start_simulation, weighted_average
from fedb200.server.fedavg import aggregate_flat

NOTEBOOK_PAGERANK = [0.083054, 0.105115, 0.113125, 0.089970, 0.133988, 0.086440, 0.086141, 0.082843, 0.101708, 0.117616]


def test_pagerank_reproduces_notebook_vector_thresholds_and_anomalies():
    t = Topology.reference10()
    bad, r, (lo, hi) = pagerank_anomalies(t.edge_weights())        # All_graphs_IMDB_dataset.ipynb:52-54
    np.testing.assert_allclose(r, NOTEBOOK_PAGERANK, atol=5e-7)
    assert abs(lo - 0.083493) < 1e-6 and abs(hi - 0.116507) < 1e-6
    assert bad == [0, 4, 7, 9]
    nx = pytest.importorskip("networkx")
    G = nx.DiGraph()
    w = t.edge_weights()
    for u in range(10):
        for v in range(10):
            if u != v:
                G.add_edge(u, v, weight=w[u, v])
    pr = nx.pagerank(G, weight="weight")
    np.testing.assert_allclose(r, [pr[i] for i in range(10)], atol=1e-6)


def test_pagerank_handles_dangling_nodes():
    w = np.array([[0, 1, 0], [0, 0, 1], [0, 0, 0.0]])
    r, _ = pagerank(w)
    assert abs(r.sum() - 1.0) < 1e-9 and r[2] > r[1] > r[0]


def test_other_detectors_match_notebook_outputs():
    t = Topology.reference10()
    deg = t.weighted_degree_undirected()
    assert modz_anomalies(deg)[0] == [8, 9]                         # :489-491
    assert dbscan_anomalies(deg)[0] == []                           # :331-333 (eps 300 >> feature scale)
    assert community_anomalies(t.edge_weights()) == []              # :665-666
    assert dbscan_anomalies(np.array([0.0, 0.1, 0.2, 50.0]), eps=1.0, min_samples=2)[0] == [3]
    assert detect(t, "pagerank") == [0, 4, 7, 9] and detect(t, "modz") == [8, 9] and detect(t, "none") == []


def test_info_passing_sync_is_sum_async_is_max_and_routing_helps():
    bw = np.array([[0, 10, 1], [10, 0, 10], [1, 10, 0.0]])          # 0->2 direct is slow, via 1 is fast
    t = Topology(bw, unit_bytes_per_s=1.0)
    direct = info_passing_time(t, 10.0, 0, routed=False)
    assert direct["sync"] == pytest.approx(1.0 + 10.0) and direct["async"] == pytest.approx(10.0)
    routed = info_passing_time(t, 10.0, 0)
    assert routed["sync"] == pytest.approx(1.0 + 2.0) and routed["async"] == pytest.approx(2.0)
    plan = optimise_route(t, 10.0, 0, compute_time=5.0)
    assert plan.routes[2] == [0, 1, 2] and plan.objective == pytest.approx(7.0)
    assert optimise_route(t, 10.0, 0, fanout=1).selected == [1]
    d, _ = shortest_paths(t.hop_seconds(10.0), 0)
    assert d.tolist() == pytest.approx([0.0, 1.0, 2.0])
    ref = Topology.reference10()
    nbytes = 0.4036 * 2 ** 30
    full = info_passing_time(ref, nbytes, 1)
    pruned = info_passing_time(ref, nbytes, 1, removed=[0, 4, 7, 9])
    assert pruned["sync"] < full["sync"] and full["async"] < 0.3 * full["sync"]     # async << sync (README: -76 %)


def test_fedavg_weighted_mean_closed_form_and_batches_as_weights():
    a = [np.full((2, 2), 1.0, np.float32), np.full(3, 10.0, np.float32)]
    b = [np.full((2, 2), 4.0, np.float32), np.full(3, 40.0, np.float32)]
    out = aggregate_weighted([(a, 1), (b, 3)])
    np.testing.assert_allclose(out[0], 3.25)
    np.testing.assert_allclose(out[1], 32.5)
    flat = aggregate_flat([torch.ones(5), torch.full((5,), 4.0)], [1, 3])
    torch.testing.assert_close(flat, torch.full((5,), 3.25))
    masked = aggregate_flat([torch.ones(5), torch.full((5,), 4.0)], [1, 3], mask=[True, False])
    torch.testing.assert_close(masked, torch.ones(5))
    m = weighted_average([(10, {"accuracy": 0.5, "loss": 1.0}), (30, {"accuracy": 0.9, "loss": 0.2})])
    assert m["accuracy"] == pytest.approx(0.8) and m["loss"] == pytest.approx(0.4)


def test_start_simulation_round_structure_with_flower_defaults():
    class C:
        def __init__(self, cid):
            self.cid = int(cid)

        def get_parameters(self, config):
            return [np.zeros(2, np.float32)]

        def fit(self, params, config):
            return [params[0] + self.cid + 1], self.cid + 1, {}

        def evaluate(self, params, config):
            return float(params[0][0]), 10, {"accuracy": 0.5, "loss": float(params[0][0])}

    hist = start_simulation(lambda cid: C(cid), 3, ServerConfig(num_rounds=2),
                            FedAvg(evaluate_metrics_aggregation_fn=weighted_average, seed=0), seed=0)
    # round 1: sum n_k (w0 + k + 1) / sum n_k with n_k = k+1 -> (1*1 + 2*2 + 3*3)/6 = 14/6
    assert hist.losses_distributed[0][1] == pytest.approx(14 / 6)
    assert hist.losses_distributed[1][1] == pytest.approx(28 / 6)
    assert hist.metrics_distributed["accuracy"] == [(1, 0.5), (2, 0.5)]
    assert FedAvg(seed=0).configure_fit(1, 1) == []           # min_available_clients = 2


def test_mixing_matrices_masking_and_unweighted_mean():
    full = mixing_matrix(np.ones((4, 4)))
    np.testing.assert_allclose(full, 0.25)
    ring = mixing_matrix(Topology.ring(5).adjacency(), "metropolis")
    np.testing.assert_allclose(ring.sum(1), 1.0)
    np.testing.assert_allclose(ring, ring.T)
    m = apply_mask(full, [2])
    assert np.all(m[:, 2] == 0) and np.allclose(m.sum(1), 1.0) and m[0, 0] == pytest.approx(1 / 3)
    flats = [torch.full((8,), float(i)) for i in range(4)]
    torch.testing.assert_close(mix_flat(flats, full[0]), torch.full((8,), 1.5))       # sum(p)/len(p)


def test_cosine_pagerank_scorer_flags_poisoned_updates_and_injector():
    torch.manual_seed(0)
    base = torch.randn(2000)
    common = torch.randn(2000)
    ups = [base + 0.01 * (common + 0.5 * torch.randn(2000)) for _ in range(8)]
    inj = AnomalyInjector(8, 2, "signflip", strength=5.0, seed=3)
    tg = inj.targets()
    assert len(tg) == 2
    for k in tg:
        assert inj.maybe_corrupt(k, ups[k], base, 0)
    g, cos = cosine_gram([u - base for u in ups])
    bad, ranks, _ = score_updates(g)
    assert sorted(bad) == tg
    good = [i for i in range(8) if i not in tg]
    assert cos[good[0], good[1]] > 0.5 and cos[good[0], tg[0]] < 0
    x = base.clone() + 0.01
    inject_anomaly(x, base, "scale", 10.0)
    torch.testing.assert_close(x - base, torch.full_like(base, 0.1), rtol=1e-4, atol=1e-6)


def test_async_engine_staleness_weights_replay_and_max_vs_sum():
    t = Topology.reference10()
    rng = np.random.default_rng(0)
    speeds = rng.uniform(0.5, 2.0, size=10)
    mixes = []
    eng = AsyncGossipEngine(t, 0.0435 * 2 ** 30, lambda k: float(speeds[k]),
                            lambda k, s, w, sw: mixes.append((k, tuple(s), tuple(w), sw)), decay=0.5, fanout=3)
    res = eng.run(4)
    assert res["mixes"] == 40 and len(mixes) == 40
    assert res["async_comm_time"] < res["sync_equiv_comm_time"]          # max over destinations < sum
    for rec, (ws, sw) in zip(eng.log, eng.replay_weights()):              # pure function of logged epochs
        assert list(ws) == pytest.approx(rec.weights) and sw == pytest.approx(rec.self_weight)
        assert abs(sum(rec.weights) + rec.self_weight - 1.0) < 1e-12 and rec.self_weight >= 0.5 - 1e-12
        # a fast peer may be AHEAD of the receiver: staleness clamps at 0, so no weight exceeds the base share
        assert all(0.0 < w <= eng.mix_rate / len(rec.sources) + 1e-12 for w in rec.weights)
    assert staleness_weight(0.4, 0) == 0.4 and staleness_weight(0.4, 3, 0.5) == pytest.approx(0.2)
    assert any(any(e < r.epoch for e in r.source_epochs) for r in eng.log)        # stale updates really occur
    eng2 = AsyncGossipEngine(t, 1e6, lambda k: 1.0, lambda *a: None, masked=[0, 4])
    eng2.run(2)
    assert all(r.peer not in (0, 4) and 0 not in r.sources and 4 not in r.sources for r in eng2.log)


def test_merkle_root_rules_against_hashlib():
    data = bytes(range(256)) * 10
    leaves = leaf_digests(data, 256)
    assert len(leaves) == 10 and leaves[0] == hashlib.sha256(data[:256] + b"\x00").digest()
    lvl = reduce_level(leaves[:3])
    assert lvl[0] == hashlib.sha256(leaves[0] + leaves[1] + b"\x01").digest() and lvl[1] == leaves[2]      # promotion
    assert merkle_root(data[:256], 256) == hashlib.sha256(data[:256] + b"\x00").digest()
    assert merkle_root(data, 256) != merkle_root(data[:-1] + b"\x00", 256)
    # domain separation: the 64-byte concatenation of two leaf digests, presented as ONE leaf, does not hash to their parent
    assert merkle_root(leaves[0] + leaves[1], 64) != reduce_level(leaves[:2])[0]


def test_ledger_chain_validation_and_tamper_detection(tmp_path):
    led = Ledger(chunk_bytes=64, clock=lambda: 42)
    xs = [np.arange(100, dtype=np.float32) * (i + 1) for i in range(5)]
    for i, x in enumerate(xs):
        led.record_update(1, i, (i + 1) % 5, x.tobytes(), sample_count=4, route=[i, 2, (i + 1) % 5])
    assert led.height == 5 and led.validate() == -1
    assert led.verify_update(2, xs[2].tobytes()) and not led.verify_update(2, xs[3].tobytes())
    assert BlockHeader.unpack(led.blocks[3].pack()) == led.blocks[3] and len(led.blocks[3].pack()) == 128
    led.save(str(tmp_path / "l.json"))
    led2 = Ledger.load(str(tmp_path / "l.json"))
    assert led2.validate() == -1 and led2.head == led.head
    led2.blocks[1].sample_count = 99
    assert led2.validate() == 1
    with pytest.raises(LedgerError):
        led2.assert_valid()
    led3 = Ledger.load(str(tmp_path / "l.json"))
    led3.blocks[2].prev_hash = b"\x01" * 32
    assert led3.validate() == 2
    pow_led = Ledger(chunk_bytes=64, difficulty_bits=8)
    blk = pow_led.record_update(0, 0, 1, xs[0].tobytes())
    assert blk.hash()[0] == 0 and pow_led.validate() == -1


def test_report_tables_match_detectors_and_timing_model(tmp_path):
    from fedb200.utils import info_passing_table, sweep_table, write_csv
    rows = info_passing_table()
    by = {r["detector"]: r for r in rows}
    assert by["pagerank"]["removed"] == [0, 4, 7, 9] and by["modz"]["removed"] == [8, 9] and by["dbscan"]["removed"] == []
    assert all(r["async_s"] <= r["sync_s"] for r in rows)
    assert by["pagerank"]["sync_s"] < by["none"]["sync_s"]          # fewer destinations after elimination
    assert all(r["async_saving_pct"] > 50 for r in rows)            # async = max << sync = sum (README: -76 %)
    write_csv(rows, str(tmp_path / "t.csv"))
    assert (tmp_path / "t.csv").read_text().startswith("detector,removed,source,sync_s,async_s")
    sw = sweep_table({5: {"latency_min": 1.0, "global_accuracies": [0.5, 0.6]}, 10: {"latency_min": 2.0, "global_accuracies": [0.7]}})
    assert [r["workers"] for r in sw] == [5, 10] and sw[0]["final_accuracy"] == 0.6


def test_result_figures_svg(tmp_path):
    """C26: the notebooks' figure set is written as SVG from computed tables (no matplotlib in the image)."""
    import json
    import xml.etree.ElementTree as ET
    from fedb200.utils import plots
    sweeps = {"centralised": {5: {"latency_min": 1.0, "global_accuracies": [0.5, 0.8], "memory_growth_gb": 0.2},
                              10: {"latency_min": 2.0, "global_accuracies": [0.6, 0.85], "memory_growth_gb": 0.3}},
              "decentralised": {5: {"latency_min": 0.7, "global_accuracies": [0.55, 0.82], "memory_growth_gb": 0.1}}}
    curves = {"serverless-IID": {"global_accuracies": [0.5, 0.7, 0.8]}, "server-IID": {"global_accuracies": [0.5, 0.6, 0.75]}}
    (tmp_path / "s.json").write_text(json.dumps(sweeps))
    (tmp_path / "c.json").write_text(json.dumps(curves))
    written = plots.main(["--out", str(tmp_path / "fig"), "--results", str(tmp_path / "s.json"), "--curves", str(tmp_path / "c.json")])
    assert len(written) == 5
    for p in written:
        root = ET.parse(p).getroot()                      # well-formed XML
        assert root.tag.endswith("svg")
    info = ET.parse(written[0]).getroot()
    bars = [e for e in info.iter() if e.tag.endswith("rect") and e.find("{http://www.w3.org/2000/svg}title") is not None]
    assert len(bars) == 8                                 # 4 detector settings x (sync, async)
    lines = [e for e in ET.parse(written[-1]).getroot().iter() if e.tag.endswith("polyline")]
    assert len(lines) == 2


def test_async_schedule_protocol_check():
    """protocol-level check of the pub/done flag handshake — all waits point backwards in the global order and,
    under random interleavings of the ranks, every mix reads exactly the version it was scheduled to read."""
    import numpy as np
    from fedb200.p2p import Topology
    from fedb200.p2p.schedule import build_async_schedule, model_durations, verify_schedule
    cases = [(Topology.reference10(), 10, 5, None, 20.0), (Topology.full(8), 8, 8, None, 0.05), (Topology.full(16), 16, 8, 3, 0.02),
             (Topology.ring(6), 6, 3, None, 0.2)]
    for topo, n, world, fanout, base in cases:
        for seed in range(3):
            sched = build_async_schedule(topo, 433e6, 4, model_durations(n, base_s=base, seed=seed), fanout=fanout)
            out = verify_schedule(sched, n, world, trials=8, seed=seed)
            assert out["events"] == 4 * n - 4 * len(sched.masked) and out["reads_checked"] > 0
    # a broken schedule must be caught: drop the overwrite waits of a slot-reusing publish
    sched = build_async_schedule(Topology.full(8), 433e6, 6, model_durations(8, seed=1))
    victims = [ev for ev in sched.events if ev.overwrite_waits]
    if victims:
        import pytest
        for ev in victims:
            ev.overwrite_waits = []
        with pytest.raises(AssertionError):
            verify_schedule(sched, 8, 8, trials=64, seed=0)


def test_async_visibility_lag_breaks_the_dependency_chain():
    """With ~1 ms modelled links and ~50 ms epochs every mix depends on the publish just before it (the execution
    serialises across GPUs); a visibility lag of one epoch moves every dependency at least a GPU-round back while the
    schedule stays valid under the protocol checker."""
    from fedb200.p2p import Topology
    from fedb200.p2p.schedule import build_async_schedule, model_durations, verify_schedule
    n, world = 32, 8
    dur = model_durations(n, base_s=0.05, seed=0)
    topo = Topology.full(n)

    def min_gap(s):
        idx = {(e.peer, e.epoch): e.index for e in s.events}
        gaps = [e.index - idx[(p, ep)] for e in s.events for p, ep in zip(e.sources, e.source_epochs)]
        return min(gaps), len(gaps)

    tight = build_async_schedule(topo, 1.34e9, 3, dur)
    lagged = build_async_schedule(topo, 1.34e9, 3, dur, min_latency=float(dur.min()))
    g0, n0 = min_gap(tight)
    g1, n1 = min_gap(lagged)
    assert g0 == 1 and n0 > 0                       # chain: some mix reads the version published by the previous event
    assert g1 >= world and n1 > 0                   # lag: nothing more recent than a full sweep of the ranks
    verify_schedule(lagged, n, world, trials=4)
    assert lagged.slots_per_peer <= 4


def test_balanced_placement_equalises_token_load_and_respects_slots():
    from fedb200.parallel import Placement
    rng = np.random.default_rng(3)
    cost = rng.lognormal(10, 0.8, size=32)
    blk, bal = Placement.block(32, 8), Placement.balanced(32, 8, cost)
    assert sorted(sum((bal.local_ids(r) for r in range(8)), [])) == list(range(32))
    assert all(len(bal.local_ids(r)) == 4 for r in range(8))                      # never more than P peers per GPU
    for k in range(32):
        r, l = bal.owner(k)
        assert bal.local_ids(r)[l] == k
        assert blk.owner(k) == (k // 4, k % 4)
    # the straggler shrinks to (about) the largest single client plus the smallest ones, the other ranks even out
    assert max(bal.loads(cost)) < max(blk.loads(cost)) and max(bal.loads(cost)) < 1.1 * max(cost)
    assert np.std(sorted(bal.loads(cost))[:-1]) < 0.5 * np.std(sorted(blk.loads(cost))[:-1])


def test_relay_plan_executes_two_hop_routes_through_a_third_gpu():
    from fedb200.p2p import RelayPlan
    from fedb200.parallel import Placement
    pl = Placement.block(8, 4)                                   # peers 0-1 on rank 0, 2-3 on rank 1, ...
    W = np.zeros((8, 8))
    for i in range(8):
        W[i, i], W[i, (i + 1) % 8], W[i, (i + 3) % 8] = 0.5, 0.25, 0.25
    relays = {(0, 3): 5, (1, 4): 5, (2, 5): 3, (6, 7): 0, (4, 5): 6}
    plan = RelayPlan(W, pl.owner, relays, max_slots=1)
    # 3 -> 5 -> 0: relay peer 5 lives on rank 2, neither the source's (1) nor the destination's (0) GPU -> executed
    assert plan.via[(0, 3)] == (2, 0) and plan.executed[(0, 3)] == [3, 5, 0] and plan.copies[2] == [(3, 0)]
    # 4 -> 5 -> 1: the relay shares the source's GPU (rank 2) -> direct
    assert (1, 4) not in plan.via and plan.executed[(1, 4)] == [4, 1]
    # 5 -> 3 -> 2: the relay shares the destination's GPU -> direct;  7 -> 0 -> 6: the relay's GPU is also the ... source 7 on rank 3, dst 6 on rank 3
    assert (2, 5) not in plan.via and (6, 7) not in plan.via
    # W[4, 5] is a real edge (4 hears from 5) but both live on rank 2 -> never relayed
    assert plan.executed[(4, 5)] == [5, 4]
    # every nonzero off-diagonal edge has an executed route that starts at the source and ends at the destination
    for (i, j), r in plan.executed.items():
        assert r[0] == j and r[-1] == i and W[i, j] != 0.0
    # slot budget: a second distinct source through rank 2 does not fit with max_slots = 1 -> direct
    plan2 = RelayPlan(W, pl.owner, {(0, 3): 5, (7, 0): 4}, max_slots=1)
    assert plan2.via == {(0, 3): (2, 0)} and plan2.executed[(7, 0)] == [0, 7]
    assert not RelayPlan(W, pl.owner, {}, 4).any
