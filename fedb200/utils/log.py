"""Structured JSONL metrics (replaces the reference's ``print`` calls,
src/Serverlesscase/serverless_IID_IMDB.py:265-303, and its matplotlib plots, C26)."""
from __future__ import annotations

import json
import sys
import time
from typing import Optional


class MetricsLogger:
    def __init__(self, path: Optional[str] = None, echo: bool = True, rank: int = 0):
        self.path, self.echo, self.rank = path, echo, rank
        self.records = []
        self._fh = open(path, "a") if (path and rank == 0) else None

    def log(self, event: str, **fields) -> dict:
        rec = {"event": event, "ts": time.time(), **fields}
        self.records.append(rec)
        if self.rank == 0:
            line = json.dumps(rec, default=lambda o: float(o) if hasattr(o, "__float__") else str(o))
            if self._fh:
                self._fh.write(line + "\n")
                self._fh.flush()
            if self.echo:
                print(line, file=sys.stderr)
        return rec

    def close(self) -> None:
        if self._fh:
            self._fh.close()
            self._fh = None
