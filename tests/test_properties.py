"""Property-based checks (hypothesis) of the host-side invariants the device kernels are tested against."""
import hashlib

import numpy as np
import pytest

pytest.importorskip("hypothesis")
from hypothesis import given, settings, strategies as st

from fedb200.ledger.merkle import LEAF_TAG, NODE_TAG, leaf_digests, merkle_root, reduce_level, root_from_leaves
from fedb200.p2p.gossip import apply_mask, mixing_matrix
from fedb200.p2p.observed import stale_table
from fedb200.parallel.placement import Placement, peers_per_rank


@settings(max_examples=60, deadline=None)
@given(st.binary(min_size=1, max_size=700), st.sampled_from([16, 64, 100, 4096]), st.data())
def test_merkle_root_changes_with_any_bit_and_matches_leaf_reduction(data, chunk, draw):
    root = merkle_root(data, chunk)
    leaves = np.frombuffer(b"".join(leaf_digests(data, chunk)), dtype=np.uint8).reshape(-1, 32)
    assert root_from_leaves(leaves) == root
    i = draw.draw(st.integers(0, len(data) * 8 - 1))
    flipped = bytearray(data)
    flipped[i // 8] ^= 1 << (i % 8)
    assert merkle_root(bytes(flipped), chunk) != root
    assert merkle_root(data + b"\x00", chunk) != root                  # length is part of the message (no zero-padding collisions)


@settings(max_examples=40, deadline=None)
@given(st.binary(min_size=33, max_size=64), st.binary(min_size=1, max_size=64))
def test_merkle_domain_separation_interior_pair_is_not_a_leaf(a, b):
    """The classic second-preimage trick — presenting the concatenated child digests as one leaf — must give another root."""
    la, lb = hashlib.sha256(a + LEAF_TAG).digest(), hashlib.sha256(b + LEAF_TAG).digest()
    two_leaf_root = reduce_level([la, lb])[0]
    assert two_leaf_root == hashlib.sha256(la + lb + NODE_TAG).digest()
    forged_leaf_root = merkle_root(la + lb, 64)                         # one 64-byte "leaf" holding the two digests
    assert forged_leaf_root != two_leaf_root


@settings(max_examples=60, deadline=None)
@given(st.integers(1, 8), st.integers(1, 40), st.data())
def test_balanced_placement_is_a_bijection_within_slots_and_beats_or_ties_block(world, n, draw):
    cost = draw.draw(st.lists(st.floats(1.0, 1000.0), min_size=n, max_size=n))
    P = peers_per_rank(n, world)
    bal, blk = Placement.balanced(n, world, cost), Placement.block(n, world)
    seen = set()
    for r in range(world):
        ids = bal.local_ids(r)
        assert len(ids) <= P
        for slot, peer in enumerate(ids):
            assert bal.owner(peer) == (r, slot)
            seen.add(peer)
    assert seen == set(range(n))
    # LPT under a slot cap: never worse than the contiguous-block placement by more than the largest single cost
    assert max(bal.loads(cost)) <= max(blk.loads(cost)) + max(cost) + 1e-9
    assert abs(sum(bal.loads(cost)) - sum(cost)) < 1e-6 * max(1.0, sum(cost))


@settings(max_examples=60, deadline=None)
@given(st.integers(2, 9), st.data())
def test_mixing_matrices_are_row_stochastic_and_masking_removes_columns(n, draw):
    adj = np.array(draw.draw(st.lists(st.lists(st.integers(0, 1), min_size=n, max_size=n), min_size=n, max_size=n)))
    sym = ((adj + adj.T) > 0).astype(int)
    for kind, a in (("uniform", adj), ("metropolis", sym)):
        w = mixing_matrix(a, kind)
        assert np.allclose(w.sum(axis=1), 1.0) and (w >= -1e-12).all()
        if kind == "metropolis":
            assert np.allclose(w, w.T)                                  # doubly stochastic on an undirected graph
        masked = draw.draw(st.sets(st.integers(0, n - 1), max_size=n - 1))
        wm = apply_mask(w, masked)
        assert np.allclose(wm.sum(axis=1), 1.0)
        for m in masked:
            others = [i for i in range(n) if i != m]
            assert np.allclose(wm[others, m], 0.0)                      # nobody hears from a masked peer


@settings(max_examples=40, deadline=None)
@given(st.floats(0.05, 1.0), st.floats(0.0, 3.0))
def test_staleness_table_is_positive_bounded_and_non_increasing(rate, decay):
    t = stale_table(rate, decay)
    assert len(t) == 32 and all(x > 0 for x in t) and t[0] <= rate * (1 + 1e-6)
    assert all(t[i + 1] <= t[i] for i in range(31))
