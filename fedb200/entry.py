"""Shared ``main`` for the four README entrypoints (README.md:2-5 names ``server_IID.py``,
``server_NonIID.py``, ``serverless_IID.py``, ``serverless_NonIID.py``; in the reference they are
the per-dataset scripts under ``src/Servercase`` and ``src/Serverlesscase``)."""
from __future__ import annotations

import json
import sys

from .config import FedConfig, build_parser, config_from_args
from .runtime import run_client_sweep, run_experiment


def main(mode: str, split: str, description: str, argv=None) -> dict:
    d = FedConfig(mode=mode)
    d.data.split = split
    if split == "shard":
        d.data.train_per_client, d.data.test_per_client = 240, 60     # serverless_NonIID_IMDB.py:59-60
    p = build_parser(d, description)
    p.add_argument("--sweep", default=None, help="comma-separated client counts, e.g. 5,10,20")
    p.add_argument("--quiet", action="store_true")
    args = p.parse_args(argv)
    cfg = config_from_args(args, d)
    if args.sweep:
        res = run_client_sweep(cfg, [int(x) for x in args.sweep.split(",")], echo=not args.quiet)
        summary = {n: {"global_accuracies": r["global_accuracies"], "latency_min": r["latency_min"],
                       "memory_usage_gb": r["memory_usage_gb"], "cpu_overhead_pct": r["cpu_overhead_pct"],
                       "ledger_height": r["ledger_height"]} for n, r in res.items()}
    else:
        res = run_experiment(cfg, echo=not args.quiet)
        summary = {"global_accuracies": res["global_accuracies"], "latency_min": res["latency_min"],
                   "cpu_overhead_pct": res["cpu_overhead_pct"], "memory_usage_gb": res["memory_usage_gb"],
                   "ledger_height": res["ledger_height"]}
    import os
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps(summary))
    return res
