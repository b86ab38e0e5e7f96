"""The cross-GPU flag protocol (csrc/peer_flags.cuh) explored exhaustively as a state machine: safety and progress for every
interleaving, with and without co-resident CTAs; broken variants must be caught (fedb200/utils/protocol_check.py)."""
import pytest

from fedb200.utils.protocol_check import check, check_broken_variant


@pytest.mark.parametrize("ranks,ctas,launches,resident", [(2, 2, 2, 1), (2, 2, 2, 2), (2, 3, 1, 1)])
def test_flag_protocol_is_safe_and_live_for_every_interleaving(ranks, ctas, launches, resident):
    out = check(ranks, ctas, launches, resident)
    assert out["final_states"] == 1 and out["states"] > 1000


def test_checker_catches_broken_protocols():
    with pytest.raises(AssertionError, match="end safety"):
        check(2, 2, 2, 1, mutation="no_end_wait")
    with pytest.raises(AssertionError, match="weight of epoch"):
        check(2, 2, 2, 1, mutation="epoch_before_weight")
    # the round-1 start barrier (block 0 releases the other blocks) hangs when the grid is not co-resident
    assert check_broken_variant().startswith("stuck")


def test_async_ring_never_accepts_a_lapped_chunk_once_flags_are_retired_first():
    """The outbox ring of the asynchronous mixer: retiring a slot's landing flags before its bytes change makes every lapped
    read visible to the reader's post-read check; without that step the exploration finds the unnoticed torn chunk."""
    from fedb200.utils.protocol_check import check_ring
    for ring, chunks, versions in ((2, 2, 4), (3, 2, 5), (3, 3, 6)):
        out = check_ring(ring, chunks, versions, invalidate=True)
        assert out["accepted_reads"] > 0 and out["torn_reads_detected"] > 0
    with pytest.raises(AssertionError, match="undetected torn read"):
        check_ring(2, 2, 4, invalidate=False)
