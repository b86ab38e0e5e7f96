"""The allocation-free Merkle-root entry point (``fedb200_sha256_root_static``).

Who needs it: code that must hash a weight update but cannot hand the library a workspace — the asynchronous
gossip receive path (a cudaMalloc there would synchronise the device under every other peer's kernels) and
ledger auditors that run with a fixed memory budget and no allocator.  Its leaf/level workspace is a static
device array inside the library, shared by all calls and never cleared, so this test calls it twice in a row
(a large input, then a smaller one that leaves stale digests behind it in the array) through the plain C ABI
and requires bit-identical roots to the existing two-call path with caller-owned buffers.  Collected last on
purpose (``test_zz_``): a ``pytest -x`` run reaches it only after every other test has run.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _root_existing(lib, x: torch.Tensor, chunk: int) -> bytes:
    """Existing entry points: fedb200_sha256_leaves + fedb200_merkle_root with buffers owned by the caller."""
    nbytes = x.numel() * x.element_size()
    n = (nbytes + chunk - 1) // chunk
    leaves = torch.empty(n, 32, dtype=torch.uint8, device=x.device)
    scratch = torch.empty((n + 1) // 2 + 1, 32, dtype=torch.uint8, device=x.device)
    root = torch.empty(32, dtype=torch.uint8, device=x.device)
    stream = torch.cuda.current_stream().cuda_stream
    assert lib.fedb200_sha256_leaves(x.data_ptr(), None, nbytes, chunk, leaves.data_ptr(), stream) == 0
    assert lib.fedb200_merkle_root(leaves.data_ptr(), scratch.data_ptr(), n, root.data_ptr(), stream) == 0
    return bytes(root.cpu().numpy())


def _root_static(lib, x: torch.Tensor, chunk: int, root: torch.Tensor) -> int:
    nbytes = x.numel() * x.element_size()
    return lib.fedb200_sha256_root_static(x.data_ptr(), nbytes, chunk, root.data_ptr(),
                                          torch.cuda.current_stream().cuda_stream)


def test_static_merkle_root_matches_existing_path_twice_in_a_row(cuda_lib):
    from fedb200.ledger.merkle import merkle_root
    lib = cuda_lib
    torch.manual_seed(123)
    chunk = 4096
    big = torch.randn(5_000_003, device="cuda")          # 4 883 leaves, ragged last chunk
    small = torch.randn(70_001, device="cuda")           # 69 leaves: stale digests of `big` stay behind them
    root = torch.empty(32, dtype=torch.uint8, device="cuda")   # output buffer: allocated before, not by, the calls
    torch.cuda.synchronize()
    for x in (big, small):
        want = _root_existing(lib, x, chunk)
        rc = _root_static(lib, x, chunk, root)
        assert rc == 0, lib.fedb200_last_error().decode()
        got = bytes(root.cpu().numpy())
        assert got == want, "static-workspace root differs from the caller-workspace root"
        assert got == merkle_root(x.cpu().numpy().tobytes(), chunk)     # and both equal hashlib


def test_static_merkle_root_rejects_oversized_input(cuda_lib):
    lib = cuda_lib
    root = torch.empty(32, dtype=torch.uint8, device="cuda")
    x = torch.zeros(64 * 131073 // 4, device="cuda")      # 131 073 chunks of 64 bytes: one more than the workspace holds
    rc = _root_static(lib, x, 64, root)
    assert rc != 0
    msg = lib.fedb200_last_error().decode()
    assert "exceed the static workspace" in msg, msg
    assert _root_static(lib, x[: 64 * 131072 // 4], 64, root) == 0      # exactly at the limit is accepted
    torch.cuda.synchronize()
