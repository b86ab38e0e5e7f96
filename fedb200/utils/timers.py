"""Phase timers: CUDA events on the launching stream when on a GPU, ``perf_counter`` otherwise.
Replaces the reference's single ``time.time()`` pair around the whole script
(src/Servercase/server_IID_IMDB.py:63, :221)."""
from __future__ import annotations

import time
from collections import defaultdict
from contextlib import contextmanager
from typing import Callable, Dict, List

import torch


def cuda_time_ms(fn: Callable[[], None], warmup: int = 3, iters: int = 10, flush_l2: bool = True) -> float:
    """Median device time of ``fn`` in ms (CUDA events, sync on both sides, optional L2 flush)."""
    assert torch.cuda.is_available()
    scratch = torch.empty(160 * 1024 * 1024, dtype=torch.uint8, device="cuda") if flush_l2 else None
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    times = []
    for _ in range(iters):
        if scratch is not None:
            scratch.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    times.sort()
    return times[len(times) // 2]


class PhaseTimer:
    """Accumulates per-phase wall and device milliseconds; device events are resolved lazily so
    timing never forces a sync inside a round."""

    def __init__(self, use_cuda: bool = False):
        self.use_cuda = use_cuda and torch.cuda.is_available()
        self.wall: Dict[str, float] = defaultdict(float)
        self._pending: Dict[str, List] = defaultdict(list)

    @contextmanager
    def phase(self, name: str):
        t0 = time.perf_counter()
        if self.use_cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            torch.cuda.nvtx.range_push(name)          # NVTX range: phases show up in nsys / ncu timelines
        try:
            yield
        finally:
            if self.use_cuda:
                torch.cuda.nvtx.range_pop()
                b.record()
                self._pending[name].append((a, b))
            self.wall[name] += (time.perf_counter() - t0) * 1e3

    def device_ms(self) -> Dict[str, float]:
        if not self.use_cuda:
            return dict(self.wall)
        torch.cuda.synchronize()
        return {k: sum(a.elapsed_time(b) for a, b in v) for k, v in self._pending.items()}

    def reset(self) -> None:
        self.wall.clear()
        self._pending.clear()
