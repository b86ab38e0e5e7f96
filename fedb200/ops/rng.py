"""Host replica of the library-wide counter RNG used for dropout (``csrc/common.cuh``).

One 32-bit hash of ``(seed, row, col // 2)`` yields two 16-bit lanes; element ``(row, col)`` is kept iff
lane ``col % 2`` is below ``round((1 - p) * 65536)``.  Tests use this to build the exact mask a kernel
applied, so dropout paths are checked against an fp32 reference instead of only statistically."""
from __future__ import annotations

import numpy as np

_U32 = np.uint64(0xFFFFFFFF)


def _mul32(a: np.ndarray, b: int) -> np.ndarray:
    return (a.astype(np.uint64) * np.uint64(b)) & _U32


def hash32(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64) & _U32
    x ^= x >> np.uint64(16)
    x = _mul32(x, 0x7FEB352D)
    x ^= x >> np.uint64(15)
    x = _mul32(x, 0x846CA68B)
    x ^= x >> np.uint64(16)
    return x


def dropout_key(seed: int) -> int:
    return (seed & 0xFFFFFFFF) ^ ((((seed >> 32) & 0xFFFFFFFF) * 0xC2B2AE3D) & 0xFFFFFFFF)


def thresh16(p: float) -> int:
    keep = np.float32(1.0) - np.float32(p)
    return 65536 if keep >= 1.0 else int(np.float32(keep * np.float32(65536.0) + np.float32(0.5)))


def dropout_keep(seed: int, rows: np.ndarray, cols: np.ndarray, p: float) -> np.ndarray:
    """Boolean keep-mask for the broadcast of ``rows`` x ``cols`` (uint32 indices)."""
    r = np.asarray(rows, dtype=np.uint64)
    c = np.asarray(cols, dtype=np.uint64)
    word = hash32(np.uint64(dropout_key(seed)) ^ _mul32(r, 0x9E3779B1) ^ _mul32(c >> np.uint64(1), 0x85EBCA77))
    lane = (word >> (np.uint64(16) * (c & np.uint64(1)))) & np.uint64(0xFFFF)
    return lane < np.uint64(thresh16(p))


def dropout_scale(seed: int, n_rows: int, n_cols: int, p: float, row0: int = 0, col0: int = 0) -> np.ndarray:
    """``[n_rows, n_cols]`` float32 multiplier matrix (0 or 1/(1-p))."""
    rows = (np.arange(n_rows, dtype=np.uint64) + np.uint64(row0))[:, None]
    cols = (np.arange(n_cols, dtype=np.uint64) + np.uint64(col0))[None, :]
    keep = dropout_keep(seed, rows, cols, p)
    return keep.astype(np.float32) / np.float32(1.0 - p)


# ---- attention-probability dropout: one full hash per (row, 32-column chunk), a multiply-xorshift per column pair -------
# (csrc/common.cuh: attn_drop_base / attn_drop_word).  The attention forward is instruction-issue bound and the two-multiply
# hash per column pair was a third of its instructions; the chunk base keeps rows / chunks independent, the cheap per-pair
# step only has to decorrelate 16 Weyl-spaced inputs.
def attn_drop_base(seed: int, rows: np.ndarray, chunks: np.ndarray) -> np.ndarray:
    r = np.asarray(rows, dtype=np.uint64)
    c = np.asarray(chunks, dtype=np.uint64)
    return hash32(np.uint64(dropout_key(seed)) ^ _mul32(r, 0x9E3779B1) ^ _mul32(c, 0x85EBCA77))


def attn_drop_word(base: np.ndarray, pair: np.ndarray) -> np.ndarray:
    x = (base.astype(np.uint64) + _mul32(np.asarray(pair, dtype=np.uint64), 0x9E3779B9)) & _U32
    x ^= x >> np.uint64(16)
    x = _mul32(x, 0x846CA68B)
    x ^= x >> np.uint64(15)
    return x


def attn_dropout_keep(seed: int, rows: np.ndarray, cols: np.ndarray, p: float) -> np.ndarray:
    c = np.asarray(cols, dtype=np.uint64)
    word = attn_drop_word(attn_drop_base(seed, rows, c >> np.uint64(5)), (c & np.uint64(31)) >> np.uint64(1))
    lane = (word >> (np.uint64(16) * (c & np.uint64(1)))) & np.uint64(0xFFFF)
    return lane < np.uint64(thresh16(p))


def attn_dropout_scale(seed: int, n_rows: int, n_cols: int, p: float, row0: int = 0, col0: int = 0) -> np.ndarray:
    rows = (np.arange(n_rows, dtype=np.uint64) + np.uint64(row0))[:, None]
    cols = (np.arange(n_cols, dtype=np.uint64) + np.uint64(col0))[None, :]
    return attn_dropout_keep(seed, rows, cols, p).astype(np.float32) / np.float32(1.0 - p)
