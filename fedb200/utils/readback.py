"""One device→host synchronisation per round.

Every ``.item()`` on a device scalar drains the stream; a round used to do that four to six times (last training loss,
Merkle root of each published update, local and global evaluation metrics), and after each drain the GPU idles while the
host ramps the next phase up.  ``HostReadback`` queues non-blocking copies of small device tensors into one pinned buffer
and resolves all of them with a single stream synchronisation at the end of the round (the reference reads every value
eagerly: ``loss.item()`` per batch, serverless_IID_IMDB.py:118-133)."""
from __future__ import annotations

from typing import List, Tuple

import torch


class HostReadback:
    def __init__(self, device, capacity: int = 1 << 16):
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.buf = torch.empty(capacity, dtype=torch.uint8, pin_memory=self.cuda)
        self.off = 0
        self.bytes_read = 0

    def push(self, t: torch.Tensor) -> Tuple[int, int, torch.dtype, torch.Size]:
        """Queue ``t`` (any small tensor); returns a handle for :meth:`get` after :meth:`sync`."""
        flat = t.detach().contiguous().reshape(-1).view(torch.uint8)
        n = flat.numel()
        if self.off + n > self.buf.numel():
            raise RuntimeError("HostReadback buffer exhausted; call reset() once per round")
        self.buf[self.off:self.off + n].copy_(flat, non_blocking=True)
        h = (self.off, n, t.dtype, t.shape)
        self.off += (n + 15) // 16 * 16
        self.bytes_read += n
        return h

    def sync(self) -> None:
        if self.cuda:
            torch.cuda.current_stream(self.device).synchronize()

    def get(self, h) -> torch.Tensor:
        off, n, dtype, shape = h
        return self.buf[off:off + n].clone().view(dtype).reshape(shape)

    def reset(self) -> None:
        self.off = 0
