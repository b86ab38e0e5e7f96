"""Host model of the attention forward's softmax bookkeeping (csrc/attention.cu: attn_fwd4_kernel), for one (sequence, head).

The kernel never knows a row's maximum before it exponentiates a kv tile: it uses a REFERENCE value per row — a guess (0) at
first, afterwards whatever the last re-sweep set — collects the tile's true maximum in the same pass, and repeats the pass
against the true maximum only if that one lies more than ``2**8`` above the reference (or, on the first tile, more than
``2**8`` below the guess).  The running output and row sum are rescaled on that path only.  This file restates exactly that
in NumPy (fp32 accumulators, bf16-rounded probabilities, base-2 exponent with the 1/(1-p) dropout scale folded in) so the rule
can be tested on a CPU against a plain softmax; the device kernel is tested against fp32 PyTorch in tests/test_gpu_model.py.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

LOG2E = 1.4426950408889634
RESCALE_LOG2 = 8.0


def _bf16(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even to bfloat16, returned as float32."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32)


def lazy_softmax_attention(q: np.ndarray, k: np.ndarray, v: np.ndarray, scale: float, tile: int = 128,
                           keep: Optional[np.ndarray] = None, dropout_p: float = 0.0) -> Tuple[np.ndarray, np.ndarray, int]:
    """``q, k, v``: [len, d] float32 (already bf16-representable).  ``keep``: optional bool [len, len] dropout decisions.
    Returns (output [len, d], lse in base-2 units of the scaled scores [len], number of re-sweeps taken)."""
    n, d = q.shape
    inv_keep = 1.0 / (1.0 - dropout_p) if dropout_p > 0 else 1.0
    log2_inv_keep = np.float32(np.log2(inv_keep))
    scale_log2 = np.float32(scale * LOG2E)
    m_ref = np.zeros(n, dtype=np.float32)                 # the guess
    l = np.zeros(n, dtype=np.float32)
    o = np.zeros((n, d), dtype=np.float32)
    resweeps = 0
    for j, lo in enumerate(range(0, n, tile)):
        s = (q @ k[lo:lo + tile].T).astype(np.float32)    # raw scores of this kv tile
        mrow = s.max(axis=1) * scale_log2
        redo = (mrow > m_ref + RESCALE_LOG2) | ((j == 0) & (mrow < m_ref - RESCALE_LOG2))
        # the kernel decides per warp of 32 rows: if any row of the warp needs it, every row of the warp moves its reference
        for w0 in range(0, n, 32):
            rows = slice(w0, min(w0 + 32, n))
            if redo[rows].any():
                resweeps += 1
                m_new = mrow[rows] if j == 0 else np.maximum(m_ref[rows], mrow[rows])
                if j > 0:
                    alpha = np.exp2(m_ref[rows] - m_new).astype(np.float32)
                    l[rows] *= alpha
                    o[rows] *= alpha[:, None]
                m_ref[rows] = m_new
        p = np.exp2(s * scale_log2 - (m_ref - log2_inv_keep)[:, None]).astype(np.float32)
        l += p.sum(axis=1, dtype=np.float32)              # the row sum sees every probability (dropped ones too)
        if keep is not None:
            p = np.where(keep[:, lo:lo + tile], p, 0.0).astype(np.float32)
        o += _bf16(p) @ v[lo:lo + tile]                   # P is bf16 on the tensor core, accumulation fp32
    out = o * (inv_keep / l)[:, None]                     # l carries the folded 1/(1-p)
    lse2 = m_ref + np.log2(l) - log2_inv_keep
    return out.astype(np.float32), lse2.astype(np.float32), resweeps
