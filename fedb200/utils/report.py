"""Result tables the reference's notebooks plot (C26), as data instead of matplotlib figures.

The notebooks draw grouped bar charts from numeric literals: latency / accuracy / memory vs 5-10-20 workers for
centralised vs decentralised runs (All_graphs_IMDB_dataset.ipynb:731-915), information-passing time (synchronous vs
asynchronous) after removing each detector's anomalous nodes (:965-1071) and accuracy-vs-round curves (:1141-1144).
Here the same tables are *computed*: the information-passing table from the topology + detectors + timing model, the
worker-sweep and accuracy tables from run results (``run_client_sweep`` / the JSONL metrics)."""
from __future__ import annotations

import csv
import json
from typing import Dict, Iterable, List, Optional, Sequence

from ..anomaly import detect
from ..p2p import Topology, info_passing_time

DETECTORS = ("dbscan", "modz", "pagerank")                   # bar order of the notebooks
MODEL_BYTES = {"biobert": 0.4036 * 2 ** 30, "albert-base-v2": 0.043 * 2 ** 30}   # sizes quoted in the notebooks


def info_passing_table(topo: Optional[Topology] = None, model_bytes: float = MODEL_BYTES["biobert"], source: int = 0,
                       detectors: Sequence[str] = DETECTORS, routed: bool = True) -> List[Dict]:
    """One row per detector: removed nodes, synchronous (sum) and asynchronous (max) information-passing time in seconds,
    and the async saving (the README's "-76 %" claim is this ratio on the authors' numbers)."""
    topo = topo or Topology.reference10()
    rows = []
    for det in ("none",) + tuple(detectors):
        removed = [] if det == "none" else detect(topo, det)
        src = source if source not in removed else next(i for i in range(topo.n) if i not in removed)
        t = info_passing_time(topo, model_bytes, src, removed=removed, routed=routed)
        rows.append({"detector": det, "removed": removed, "source": src, "sync_s": t["sync"], "async_s": t["async"],
                     "async_saving_pct": 100.0 * (1.0 - t["async"] / t["sync"]) if t["sync"] > 0 else 0.0})
    return rows


def sweep_table(results: Dict[int, Dict]) -> List[Dict]:
    """Latency / accuracy / memory vs number of workers from ``run_client_sweep`` output."""
    return [{"workers": n, "latency_min": r.get("latency_min"), "final_accuracy": (r.get("global_accuracies") or [None])[-1],
             "memory_growth_gb": r.get("memory_growth_gb"), "cpu_overhead_pct": r.get("cpu_overhead_pct")}
            for n, r in sorted(results.items())]


def accuracy_curves(named_runs: Dict[str, Dict]) -> Dict[str, List[float]]:
    """Accuracy vs round for several runs, e.g. serverless-IID / serverless-NonIID / server-IID / server-NonIID."""
    return {name: list(r.get("global_accuracies", [])) for name, r in named_runs.items()}


def write_csv(rows: Iterable[Dict], path: str) -> None:
    rows = list(rows)
    if not rows:
        return
    with open(path, "w", newline="") as fh:
        w = csv.DictWriter(fh, fieldnames=list(rows[0].keys()))
        w.writeheader()
        for r in rows:
            w.writerow({k: (json.dumps(v) if isinstance(v, (list, dict)) else v) for k, v in r.items()})
