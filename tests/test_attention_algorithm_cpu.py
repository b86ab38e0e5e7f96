"""The forward attention kernel's softmax rule (reference maximum that lags the true one, exact re-sweep past 2^8, rescale of the
running output only then), restated on the host (fedb200/ops/attn_model.py) and compared with a plain softmax."""
import numpy as np
import pytest

from fedb200.ops.attn_model import LOG2E, _bf16, lazy_softmax_attention


def _ref(q, k, v, scale, keep=None, p=0.0):
    s = (q @ k.T) * scale
    pr = np.exp(s - s.max(axis=1, keepdims=True))
    pr /= pr.sum(axis=1, keepdims=True)
    lse2 = (np.log(np.exp(s - s.max(axis=1, keepdims=True)).sum(axis=1)) + s.max(axis=1)) * LOG2E
    if keep is not None:
        pr = np.where(keep, pr, 0.0) / (1.0 - p)
    return pr @ v, lse2


@pytest.mark.parametrize("n,amp,ramp", [(300, 0.5, False), (512, 3.0, False), (385, 3.0, True), (64, 6.0, True), (1, 1.0, False)])
def test_lazy_reference_softmax_equals_plain_softmax(n, amp, ramp):
    rng = np.random.default_rng(n)
    q, k, v = (_bf16(rng.standard_normal((n, 64)).astype(np.float32) * amp) for _ in range(3))
    if ramp:                                   # later keys score higher: the reference has to move from tile to tile
        k = _bf16(k * np.linspace(0.5, 2.0, n, dtype=np.float32)[:, None])
    out, lse2, resweeps = lazy_softmax_attention(q, k, v, 0.125)
    want, want_lse = _ref(q.astype(np.float64), k.astype(np.float64), v.astype(np.float64), 0.125)
    rel = np.linalg.norm(out - want) / max(np.linalg.norm(want), 1e-30)
    assert rel < 6e-3, rel                     # bf16 probabilities
    assert np.abs(lse2 - want_lse).max() < 2e-3 * max(1.0, np.abs(want_lse).max())
    if amp >= 3.0 and n > 1:
        assert resweeps > 0                    # the large-score cases really take the re-sweep path
    if amp <= 0.5:
        assert resweeps == 0                   # small scores stay within 2^8 of the initial guess: one sweep per tile


def test_lazy_reference_softmax_with_dropout_mask():
    rng = np.random.default_rng(7)
    n, p = 257, 0.1
    q, k, v = (_bf16(rng.standard_normal((n, 64)).astype(np.float32)) for _ in range(3))
    keep = rng.random((n, n)) >= p
    out, lse2, _ = lazy_softmax_attention(q, k, v, 0.125, keep=keep, dropout_p=p)
    want, want_lse = _ref(q.astype(np.float64), k.astype(np.float64), v.astype(np.float64), 0.125, keep, p)
    assert np.linalg.norm(out - want) / np.linalg.norm(want) < 6e-3
    assert np.abs(lse2 - want_lse).max() < 1e-3            # the lse is the one of the UNDROPPED softmax (the backward needs that)
