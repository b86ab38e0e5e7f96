"""CPU / memory / latency overhead meter (C15).

Reference prologue/epilogue: ``psutil.cpu_percent()`` delta, RSS delta in GiB — printed as
``(rss_start - rss_end)`` so growth shows up negative (recorded -3.30 GB,
serverless_cancer_classification_with_BioBERT.ipynb:705) — and ``Latency = (end-start)/60`` min
(src/Servercase/server_IID_IMDB.py:59-63, :221-233).  ``memory_usage_gb`` keeps that sign
convention for comparability; ``memory_growth_gb`` is the intuitive one."""
from __future__ import annotations

import time

try:
    import psutil
except ImportError:                       # pragma: no cover
    psutil = None


class OverheadMeter:
    def __init__(self):
        self.start()

    def start(self) -> None:
        self.t0 = time.time()
        self.cpu0 = psutil.cpu_percent() if psutil else 0.0
        self.rss0 = psutil.Process().memory_info().rss if psutil else 0

    def report(self) -> dict:
        cpu1 = psutil.cpu_percent() if psutil else 0.0
        rss1 = psutil.Process().memory_info().rss if psutil else 0
        return {"cpu_overhead_pct": cpu1 - self.cpu0,
                "memory_usage_gb": (self.rss0 - rss1) / (1024 ** 3),
                "memory_growth_gb": (rss1 - self.rss0) / (1024 ** 3),
                "latency_min": (time.time() - self.t0) / 60.0}
