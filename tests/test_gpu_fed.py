"""Federated kernels (SHA-256 / Merkle / block validation / PageRank / cosine Gram) vs host oracles."""
import hashlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ops():
    from fedb200.ops import cuda_ops
    return cuda_ops


@pytest.mark.parametrize("nfloats,chunk", [(16, 64), (1024 * 50 + 4, 4096), (3_000_001 * 1, 4096), (4096 * 3, 4096), (1021, 256)])
def test_sha256_leaves_and_root_bit_exact(cuda_lib, nfloats, chunk):
    from fedb200.ledger.merkle import leaf_digests, merkle_root
    ops = _ops()
    x = torch.randn(nfloats, device="cuda")
    dst = torch.zeros_like(x)
    leaves = ops.sha256_leaves(x, dst, chunk)
    raw = x.cpu().numpy().tobytes()
    want = leaf_digests(raw, chunk)
    got = leaves.cpu().numpy()
    assert got.shape[0] == len(want)
    for i in (0, len(want) // 2, len(want) - 1):
        assert bytes(got[i]) == want[i], f"leaf {i} differs"
    assert all(bytes(g) == w for g, w in zip(got, want))
    assert torch.equal(dst, x), "fused copy differs from the source"
    root = ops.merkle_root(leaves.clone())
    assert bytes(root.cpu().numpy()) == merkle_root(raw, chunk)
    assert ops.sha256_copy_root(x, None, chunk) == merkle_root(raw, chunk)


def test_block_validate_on_device(cuda_lib):
    from fedb200.ledger import Ledger
    ops = _ops()
    led = Ledger(chunk_bytes=256)
    rng = np.random.default_rng(0)
    for i in range(300):
        led.record_update(i // 10, i % 10, (i + 1) % 10, rng.standard_normal(64).astype(np.float32).tobytes(), 5, [i % 10, 3],
                          timestamp_ns=i)
    hdr = torch.from_numpy(led.headers_array()).cuda()
    rec = torch.from_numpy(np.frombuffer(b"".join(led.hashes), dtype=np.uint8).reshape(-1, 32).copy()).cuda()
    bad, hashes = ops.block_validate(hdr, rec)
    assert bad == -1
    assert bytes(hashes[17].cpu().numpy()) == led.hashes[17] == hashlib.sha256(led.blocks[17].pack()).digest()
    hdr2 = hdr.clone(); hdr2[123, 60] ^= 1            # flip one bit inside block 123's prev_hash
    assert ops.block_validate(hdr2, rec)[0] == 123
    hdr3 = hdr.clone(); hdr3[200, 20] ^= 1            # tamper with a payload field: hash 200 changes
    assert ops.block_validate(hdr3, rec)[0] == 200
    assert ops.block_validate(hdr3, None)[0] == 201   # without recorded hashes the next link breaks


def test_pagerank_matches_notebook_vector(cuda_lib):
    from fedb200.anomaly import pagerank_anomalies
    from fedb200.p2p import Topology
    ops = _ops()
    w = Topology.reference10().edge_weights()
    ranks, mask, info = ops.pagerank(torch.tensor(w, dtype=torch.float32, device="cuda"))
    want_bad, want_r, (lo, hi) = pagerank_anomalies(w)
    np.testing.assert_allclose(ranks.cpu().numpy(), want_r, rtol=0, atol=2e-7)
    assert sorted(np.flatnonzero(mask.cpu().numpy()).tolist()) == want_bad == [0, 4, 7, 9]
    assert abs(info[1].item() - lo) < 1e-6 and abs(info[2].item() - hi) < 1e-6


@pytest.mark.parametrize("N,D", [(8, 64 * 200), (10, 100_000), (64, 1_000_000), (128, 64 * 40 + 8)])
def test_gram_tcgen05(cuda_lib, N, D):
    ops = _ops()
    torch.manual_seed(0)
    d = (torch.randn(N, D, device="cuda") * 0.1).bfloat16()
    g = ops.gram(d)
    ref = d.float() @ d.float().T
    torch.cuda.synchronize()
    err = (g - ref).abs().max().item()
    assert err <= 2e-3 * ref.abs().max().item() + 1e-3, f"gram max err {err}"


@pytest.mark.parametrize("N,D", [(2, 300_000), (8, 1_000_003 // 8 * 8), (20, 64 * 1000)])
def test_gram_segment_stacked_for_few_clients(cuda_lib, N, D):
    """Few clients: rows are cut into segments stacked in the 128-row tile; same Gram, full-size TMA boxes."""
    ops = _ops()
    torch.manual_seed(1)
    lay = ops.gram_layout(D, 1, N)
    assert lay["S"] * N <= 128 and lay["S"] >= 4 and lay["D_pad"] >= D
    d = torch.zeros(N, lay["D_pad"], dtype=torch.bfloat16, device="cuda")
    d[:, :D] = (torch.randn(N, D, device="cuda") * 0.1).bfloat16()
    out = torch.zeros(N, N, device="cuda")
    ops.gram_seg([d.data_ptr()], N, lay, 0, out)
    ref = d.float() @ d.float().T
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item()
    assert err <= 2e-3 * ref.abs().max().item() + 1e-3, f"segmented gram max err {err}"


def test_gram_fused_single_launch_one_rank_three_epochs(cuda_lib):
    """The single-launch form (barriers, partial clearing and the cross-rank sum inside the kernel) on one GPU: three calls
    in a row with different data must each give the plain Gram — the partial buffers alternate by epoch parity and are
    cleared by the kernel itself, never by the caller."""
    import ctypes as C
    ops = _ops()
    lib = cuda_lib
    N, D = 8, 64 * 3000
    lay = ops.gram_layout(D, 1, N)
    flags = torch.zeros(128, dtype=torch.int32, device="cuda")
    part = torch.zeros(2 * N * N, device="cuda")
    res = torch.full((N, N), 7.0, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    arr = lambda t: C.cast((C.c_void_p * 1)(t.data_ptr()), C.c_void_p)
    for epoch in (1, 2, 3):
        torch.manual_seed(epoch)
        d = torch.zeros(N, lay["D_pad"], dtype=torch.bfloat16, device="cuda")
        d[:, :D] = (torch.randn(N, D, device="cuda") * 0.1).bfloat16()
        rc = lib.fedb200_gram_fused_seg(arr(d), 1, N, lay["seg_len"], lay["nseg_total"], 0, lay["S"], arr(part), arr(flags),
                                        res.data_ptr(), 0, epoch, 2000, 0, stream)
        assert rc == 0, lib.fedb200_last_error().decode()
        ref = d.float() @ d.float().T
        torch.cuda.synchronize()
        err = (res - ref).abs().max().item()
        assert err <= 2e-3 * ref.abs().max().item() + 1e-3, f"epoch {epoch}: fused gram max err {err}"
    # column-range form (S = 1)
    d = (torch.randn(40, 64 * 500, device="cuda") * 0.1).bfloat16()
    part2, res2 = torch.zeros(2 * 1600, device="cuda"), torch.empty(40, 40, device="cuda")
    rc = lib.fedb200_gram_fused(arr(d), 1, 40, d.shape[1], d.shape[1], 0, d.shape[1], arr(part2), arr(flags), res2.data_ptr(),
                                0, 4, 2000, 0, stream)
    assert rc == 0, lib.fedb200_last_error().decode()
    torch.cuda.synchronize()
    ref = d.float() @ d.float().T
    assert (res2 - ref).abs().max().item() <= 2e-3 * ref.abs().max().item() + 1e-3
    assert int(flags[18]) == 0                                  # no barrier timed out


def test_gram_pagerank_flags_poisoned_update(cuda_lib):
    from fedb200.anomaly import score_updates
    ops = _ops()
    torch.manual_seed(1)
    N, D = 8, 200_000
    common = torch.randn(D, device="cuda")
    deltas = torch.stack([common + 0.5 * torch.randn(D, device="cuda") for _ in range(N)])
    deltas[3] = -4.0 * deltas[3]
    deltas[6] = torch.randn(D, device="cuda") * 3
    g = ops.gram(deltas.bfloat16())
    ranks, mask, info = ops.pagerank(g, mode=1, low_only=True)
    bad_host, r_host, _ = score_updates(g.double().cpu().numpy())
    assert set(bad_host) <= set(np.flatnonzero(mask.cpu().numpy()).tolist())      # host adds the min-gap guard
    assert 3 in bad_host and 6 in bad_host
    np.testing.assert_allclose(ranks.cpu().numpy(), r_host, atol=1e-6)


def test_runtime_receiver_verification_uses_static_root(cuda_lib):
    """--verify-updates on the GPU: the consumer's check runs through fedb200_sha256_root_static (no workspace, no allocation
    in the receive path); an update altered after it was hashed fails the round."""
    from fedb200.config import FedConfig, ModelConfig
    from fedb200.runtime import FedRuntime
    cfg = FedConfig(mode="serverless", num_clients=1, num_rounds=2, device="cuda", backend="fused", comm="loopback",
                    ledger=True, verify_updates=True)
    cfg.model = ModelConfig.preset("bert-mini", num_labels=2)
    cfg.data.population, cfg.data.train_per_client, cfg.data.test_per_client = 400, 32, 32
    cfg.data.max_seq_len, cfg.data.batch_size = 128, 16
    rt = FedRuntime(cfg, echo=False)
    rt.round_serverless(0)
    orig = rt._publish

    def tampering_publish(li, lazy=False):
        root = orig(li, lazy=lazy)
        rt.slots[li][11] += 1.0
        return root
    rt._publish = tampering_publish
    with pytest.raises(RuntimeError, match="does not hash to the Merkle root"):
        rt.round_serverless(1)
    rt.close()


def test_deterministic_rounds_give_the_same_ledger_roots(cuda_lib):
    """--deterministic: two federated runs from the same seed write the same Merkle roots into their ledgers, so an auditor
    can re-derive an entry by re-running the round (the default, atomic-reduction mode gives different low bits every run)."""
    from fedb200.config import FedConfig, ModelConfig
    from fedb200.runtime import FedRuntime

    def roots():
        cfg = FedConfig(mode="serverless", num_clients=3, num_rounds=2, device="cuda", backend="fused", comm="loopback",
                        ledger=True, deterministic=True)
        cfg.model = ModelConfig.preset("bert-mini", num_labels=2)
        cfg.data.population, cfg.data.train_per_client, cfg.data.test_per_client = 600, 48, 32
        cfg.data.max_seq_len, cfg.data.batch_size = 160, 16
        cfg.lr = 1e-3
        rt = FedRuntime(cfg, echo=False)
        try:
            for r in range(2):
                rt.round_serverless(r)
            torch.cuda.synchronize()
            assert rt.ledger.validate() == -1
            return [b.update_root for b in rt.ledger.blocks], rt.runner.flat.clone()
        finally:
            rt.runner.set_deterministic(False)
            rt.close()
    (r1, w1), (r2, w2) = roots(), roots()
    assert len(r1) >= 6 and r1 == r2, [a.hex()[:12] + " / " + b.hex()[:12] for a, b in zip(r1, r2) if a != b]
    assert torch.equal(w1, w2)
    assert len(set(r1)) > 1                       # the roots are of different updates, not a constant


# ---- the fused multi-GPU kernels in their one-rank form (the box that runs `-m gpu` has a single GPU; the 2-8 GPU versions
# ---- of these checks live in tests/multigpu_worker.py) ----------------------------------------------------------------------
def _arr(xs):
    import ctypes as C
    return C.cast((C.c_void_p * max(len(xs), 1))(*[int(x) for x in xs]), C.c_void_p)


def test_fedavg_reduce_one_rank_two_shot_one_shot_and_masked(cuda_lib):
    lib = cuda_lib
    stream = torch.cuda.current_stream().cuda_stream
    flags = torch.zeros(128, dtype=torch.int32, device="cuda")
    n = 1_000_003
    x = torch.randn(n, device="cuda")
    out = torch.zeros(n, device="cuda")
    sh = torch.zeros(n, dtype=torch.bfloat16, device="cuda")
    for epoch, (w, ctas) in enumerate([(4.0, 0), (2.0, 8)], start=1):          # full grid, then the small overlapped grid
        rc = lib.fedb200_fedavg_reduce_ex(_arr([x.data_ptr()]), _arr([out.data_ptr()]), _arr([sh.data_ptr()]), _arr([flags.data_ptr()]),
                                          0, 1, n, w, 0.0, epoch, 2000, 0, ctas, stream)
        assert rc == 0, lib.fedb200_last_error().decode()
        torch.testing.assert_close(out, x / w, rtol=1e-6, atol=1e-7)
        torch.testing.assert_close(sh.float(), (x / w).bfloat16().float(), rtol=0, atol=0)
    keep = out.clone()                                                             # weight 0: every contributor masked -> untouched
    assert lib.fedb200_fedavg_reduce_ex(_arr([x.data_ptr()]), _arr([out.data_ptr()]), None, _arr([flags.data_ptr()]),
                                        0, 1, n, 0.0, 0.0, 3, 2000, 0, 0, stream) == 0
    assert torch.equal(out, keep)
    small, so = torch.randn(4099, device="cuda"), torch.zeros(4099, device="cuda")
    for epoch in (1, 2):                                                           # one-shot family counts its own epochs
        assert lib.fedb200_fedavg_reduce_ex(_arr([small.data_ptr()]), _arr([so.data_ptr()]), None, _arr([flags.data_ptr()]),
                                            0, 1, 4099, 1.0, 0.0, epoch, 2000, 1, 0, stream) == 0
        torch.testing.assert_close(so, small)
    tab = torch.arange(48, dtype=torch.int32, device="cuda").view(1, 48).clone()
    assert lib.fedb200_peer_exchange(_arr([tab.data_ptr()]), _arr([flags.data_ptr()]), 0, 1, 48 * 4, 1, 2000, stream) == 0
    torch.cuda.synchronize()
    assert torch.equal(tab.view(-1).cpu(), torch.arange(48, dtype=torch.int32)) and int(flags[18]) == 0


def test_gossip_mix_fused_rows_and_relay_copy(cuda_lib):
    """K2 in one launch: 5 sources -> 3 rows (every source read once), with a relay copy executed first (phase 1 + mid
    barrier): the row that is routed through the relay must read the relay buffer's copy."""
    lib = cuda_lib
    stream = torch.cuda.current_stream().cuda_stream
    flags = torch.zeros(128, dtype=torch.int32, device="cuda")
    n = 300_003
    torch.manual_seed(5)
    srcs = [torch.randn(n, device="cuda") for _ in range(4)]
    relay = torch.zeros(n, device="cuda")
    W = np.ascontiguousarray(np.random.default_rng(0).random((3, 5)).astype(np.float32))
    outs = [torch.empty(n, device="cuda") for _ in range(3)]
    all_src = srcs + [relay]                                     # source 4 = the relay buffer, filled by phase 1 from srcs[2]
    rc = lib.fedb200_gossip_mix(_arr([s.data_ptr() for s in all_src]), W.ctypes.data, 5, _arr([o.data_ptr() for o in outs]), 3,
                                _arr([srcs[2].data_ptr()]), _arr([relay.data_ptr()]), 1, 1, _arr([flags.data_ptr()]), 0, 1, n, 1, 2000,
                                stream)
    assert rc == 0, lib.fedb200_last_error().decode()
    torch.cuda.synchronize()
    assert torch.equal(relay, srcs[2])
    S = torch.stack(srcs + [srcs[2]])
    want = torch.tensor(W, device="cuda") @ S
    for r in range(3):
        torch.testing.assert_close(outs[r], want[r], rtol=1e-5, atol=1e-5)
    assert int(flags[18]) == 0


def test_gossip_mix_async_observed_versions_staleness_mask_and_replay(cuda_lib):
    """K3 one-rank: three 'peers' publish into their rings with per-chunk landing flags; the mix reads the pub words on the
    device, logs the versions it chose, discounts by staleness, skips a masked peer and a peer that never published."""
    from fedb200.p2p.observed import replay_mix, stale_table
    ops = _ops()
    n, R, chunk = 1_000_004, 3, 65536
    nch = (n + chunk - 1) // chunk
    torch.manual_seed(9)
    rings = [torch.zeros(R, n, device="cuda") for _ in range(4)]
    cfl = [torch.zeros(R, nch, dtype=torch.int32, device="cuda") for _ in range(4)]
    pub = torch.zeros(4, dtype=torch.int32, device="cuda")
    published = {}
    for peer, versions in ((0, (1, 2, 3, 4)), (1, (1, 2)), (2, (1,))):           # peer 3 never publishes; version 4 reuses slot 1
        for v in versions:
            w = torch.randn(n, device="cuda")
            published[(peer, v)] = w.cpu()
            pub[peer] = v
            ops.copy_flagged(w, rings[peer][v % R], chunk, cfl[peer][v % R], v)
    mask = torch.tensor([0, 0, 1, 0], dtype=torch.int32, device="cuda")             # peer 2 eliminated by the scorer
    me, out = torch.randn(n, device="cuda"), torch.empty(n, device="cuda")
    decision = torch.zeros(16, dtype=torch.int32, device="cuda")
    log = torch.zeros(8, dtype=torch.int32, device="cuda")
    err = torch.zeros(4, dtype=torch.int32, device="cuda")
    table = stale_table(0.5, 0.5)
    ops.gossip_mix_async(me, out, rings, cfl, [pub[i:i + 1] for i in range(4)], [0, 1, 2, 3], mask, decision, log, err.data_ptr(),
                         R, chunk, 5, 1, table, timeout_ms=2000)
    torch.cuda.synchronize()
    vers = log[:4].cpu().tolist()
    assert vers == [4, 2, 0, 0] and int(log[4]) == 0 and int(err[0]) == 0          # observed versions, no torn chunk, no timeout
    want = replay_mix(me.cpu(), vers, published, [0, 1, 2, 3], 5, table)
    torch.testing.assert_close(out.cpu(), want, rtol=1e-5, atol=1e-6)
    # second launch on the same decision buffer (launch_id 2), nothing new published: same versions, larger staleness
    ops.gossip_mix_async(me, out, rings, cfl, [pub[i:i + 1] for i in range(4)], [0, 1, 2, 3], None, decision, log, err.data_ptr(),
                         R, chunk, 7, 2, table, timeout_ms=2000)
    torch.cuda.synchronize()
    assert log[:4].cpu().tolist() == [4, 2, 1, 0]
    torch.testing.assert_close(out.cpu(), replay_mix(me.cpu(), [4, 2, 1, 0], published, [0, 1, 2, 3], 7, table), rtol=1e-5, atol=1e-6)


def test_observed_async_runtime_on_one_gpu(cuda_lib):
    """The observed-version executor end to end on one GPU (4 peers time-sharing it): kernels, ledger, replay."""
    from fedb200.config import FedConfig, ModelConfig
    from fedb200.p2p.observed import replay_mix
    from fedb200.runtime import FedRuntime
    cfg = FedConfig(mode="serverless", num_clients=4, num_rounds=3, device="cuda", backend="fused", comm="loopback",
                    asynchronous=True, async_mode="observed", async_fanin=2, ledger=True, async_chunk_kb=256)
    cfg.model = ModelConfig.preset("bert-mini", num_labels=2)
    cfg.data.population, cfg.data.train_per_client, cfg.data.test_per_client = 400, 32, 32
    cfg.data.max_seq_len, cfg.data.batch_size = 128, 16
    rt = FedRuntime(cfg, echo=False)
    rt._audit_async = True
    out = rt.run()["last"]
    recs, aud, table = rt.async_records, rt.async_audit, rt.async_table
    rt.close()
    assert out["mixes"] == 12 and out["torn_chunks"] == 0 and rt.ledger.height == 12 and rt.ledger.validate() == -1
    for r in recs:
        want = replay_mix(aud["trained"][(r["peer"], r["epoch"])], r["versions"], aud["published"], r["sources"], r["epoch"], table)
        torch.testing.assert_close(aud["published"][(r["peer"], r["epoch"])], want, rtol=1e-5, atol=1e-6)
