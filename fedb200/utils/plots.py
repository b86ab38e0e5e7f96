"""Result figures of the reference's notebooks (C26) as dependency-free SVG files.

The notebooks draw, with matplotlib, grouped bar charts (latency / accuracy / memory vs 5-10-20 workers for the
centralised and decentralised runs, All_graphs_IMDB_dataset.ipynb:731-915; information-passing time, synchronous vs
asynchronous, per anomaly detector, :965-1071) and accuracy-vs-round line charts with four curves (:1141-1144).
matplotlib is not part of this image, so the same figures are written as plain SVG from the tables in
:mod:`fedb200.utils.report` — the numbers are computed by this framework, not copied literals.

    python -m fedb200.utils.plots --out figures/            # information-passing figure from the reference topology
    python -m fedb200.utils.plots --out figures/ --results sweep.json --curves runs.json
"""
from __future__ import annotations

import argparse
import json
import os
from typing import Dict, List, Optional, Sequence

PALETTE = ("#4C72B0", "#DD8452", "#55A868", "#C44E52", "#8172B3", "#937860")
W, H, ML, MR, MT, MB = 640, 400, 64, 16, 40, 56          # canvas and margins


def _esc(s: str) -> str:
    return str(s).replace("&", "&amp;").replace("<", "&lt;").replace(">", "&gt;")


def _nice_max(v: float) -> float:
    if v <= 0:
        return 1.0
    import math
    e = 10 ** math.floor(math.log10(v))
    for m in (1, 2, 2.5, 5, 10):
        if v <= m * e:
            return m * e
    return 10 * e


def _frame(title: str, ylabel: str, ymax: float, xlabel: str = "") -> List[str]:
    out = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{W}" height="{H}" viewBox="0 0 {W} {H}" font-family="sans-serif">',
           f'<rect width="{W}" height="{H}" fill="white"/>',
           f'<text x="{W / 2}" y="22" text-anchor="middle" font-size="15">{_esc(title)}</text>',
           f'<text x="16" y="{(MT + H - MB) / 2}" text-anchor="middle" font-size="12" transform="rotate(-90 16 {(MT + H - MB) / 2})">{_esc(ylabel)}</text>']
    if xlabel:
        out.append(f'<text x="{(ML + W - MR) / 2}" y="{H - 8}" text-anchor="middle" font-size="12">{_esc(xlabel)}</text>')
    for k in range(6):                                     # horizontal grid + tick labels
        y = H - MB - (H - MB - MT) * k / 5
        out.append(f'<line x1="{ML}" y1="{y:.1f}" x2="{W - MR}" y2="{y:.1f}" stroke="#ddd"/>')
        out.append(f'<text x="{ML - 6}" y="{y + 4:.1f}" text-anchor="end" font-size="11">{ymax * k / 5:g}</text>')
    out.append(f'<line x1="{ML}" y1="{H - MB}" x2="{W - MR}" y2="{H - MB}" stroke="black"/>')
    out.append(f'<line x1="{ML}" y1="{MT}" x2="{ML}" y2="{H - MB}" stroke="black"/>')
    return out


def _legend(names: Sequence[str]) -> List[str]:
    out, x = [], ML + 8
    for i, n in enumerate(names):
        out.append(f'<rect x="{x}" y="{MT - 12}" width="10" height="10" fill="{PALETTE[i % len(PALETTE)]}"/>')
        out.append(f'<text x="{x + 14}" y="{MT - 3}" font-size="11">{_esc(n)}</text>')
        x += 22 + 7 * len(str(n))
    return out


def bar_chart_svg(groups: Sequence[str], series: Dict[str, Sequence[Optional[float]]], title: str, ylabel: str,
                  path: Optional[str] = None, xlabel: str = "") -> str:
    """Grouped bars: one group per entry of ``groups``, one bar per series (``series[name][g]``; None = no bar)."""
    vals = [v for s in series.values() for v in s if v is not None]
    ymax = _nice_max(max(vals) if vals else 1.0)
    out = _frame(title, ylabel, ymax, xlabel) + _legend(list(series))
    gw = (W - ML - MR) / max(1, len(groups))
    bw = gw * 0.8 / max(1, len(series))
    for g, gname in enumerate(groups):
        x0 = ML + g * gw + gw * 0.1
        for i, (name, s) in enumerate(series.items()):
            v = s[g] if g < len(s) else None
            if v is None:
                continue
            h = (H - MB - MT) * max(0.0, float(v)) / ymax
            out.append(f'<rect x="{x0 + i * bw:.1f}" y="{H - MB - h:.1f}" width="{bw * 0.92:.1f}" height="{h:.1f}" '
                       f'fill="{PALETTE[i % len(PALETTE)]}"><title>{_esc(name)} / {_esc(gname)}: {v:g}</title></rect>')
        out.append(f'<text x="{ML + (g + 0.5) * gw:.1f}" y="{H - MB + 16}" text-anchor="middle" font-size="11">{_esc(gname)}</text>')
    svg = "\n".join(out + ["</svg>"])
    if path:
        with open(path, "w") as fh:
            fh.write(svg)
    return svg


def line_chart_svg(curves: Dict[str, Sequence[float]], title: str, ylabel: str, path: Optional[str] = None,
                   xlabel: str = "round") -> str:
    """One polyline per named curve (accuracy vs round in the notebooks)."""
    n = max((len(c) for c in curves.values()), default=1)
    vals = [v for c in curves.values() for v in c]
    ymax = _nice_max(max(vals) if vals else 1.0)
    out = _frame(title, ylabel, ymax, xlabel) + _legend(list(curves))
    xs = lambda k: ML + (W - ML - MR) * (k / max(1, n - 1))
    for k in range(n):
        out.append(f'<text x="{xs(k):.1f}" y="{H - MB + 16}" text-anchor="middle" font-size="11">{k + 1}</text>')
    for i, (name, c) in enumerate(curves.items()):
        pts = " ".join(f"{xs(k):.1f},{H - MB - (H - MB - MT) * float(v) / ymax:.1f}" for k, v in enumerate(c))
        col = PALETTE[i % len(PALETTE)]
        out.append(f'<polyline fill="none" stroke="{col}" stroke-width="2" points="{pts}"><title>{_esc(name)}</title></polyline>')
        for k, v in enumerate(c):
            out.append(f'<circle cx="{xs(k):.1f}" cy="{H - MB - (H - MB - MT) * float(v) / ymax:.1f}" r="3" fill="{col}"/>')
    svg = "\n".join(out + ["</svg>"])
    if path:
        with open(path, "w") as fh:
            fh.write(svg)
    return svg


# ---- the notebooks' figure set -------------------------------------------------------------------------------------
def info_passing_figure(path: Optional[str] = None, **kw) -> str:
    """Information-passing time, synchronous vs asynchronous, after removing each detector's anomalous nodes."""
    from .report import info_passing_table
    rows = info_passing_table(**kw)
    return bar_chart_svg([r["detector"] for r in rows],
                         {"synchronous": [r["sync_s"] for r in rows], "asynchronous": [r["async_s"] for r in rows]},
                         "Information passing time by anomaly detector", "seconds", path, "nodes removed by")


def sweep_figures(named_sweeps: Dict[str, Dict], out_dir: str) -> List[str]:
    """Latency / accuracy / memory vs number of workers; ``named_sweeps`` maps a run name (e.g. "centralised",
    "decentralised") to ``run_client_sweep`` output (``{workers: result}``)."""
    from .report import sweep_table
    tables = {name: sweep_table({int(k): v for k, v in res.items()}) for name, res in named_sweeps.items()}
    workers = sorted({r["workers"] for t in tables.values() for r in t})
    paths = []
    for key, title, unit in (("latency_min", "Latency vs workers", "minutes"), ("final_accuracy", "Accuracy vs workers", "accuracy"),
                             ("memory_growth_gb", "Memory overhead vs workers", "GB")):
        series = {name: [next((r[key] for r in t if r["workers"] == w), None) for w in workers] for name, t in tables.items()}
        p = os.path.join(out_dir, f"sweep_{key}.svg")
        bar_chart_svg([str(w) for w in workers], series, title, unit, p, "workers")
        paths.append(p)
    return paths


def main(argv: Optional[Sequence[str]] = None) -> List[str]:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--out", default="figures")
    ap.add_argument("--results", default=None, help="JSON: {run name: {workers: run_client_sweep result}}")
    ap.add_argument("--curves", default=None, help="JSON: {run name: {'global_accuracies': [...]}}")
    a = ap.parse_args(argv)
    os.makedirs(a.out, exist_ok=True)
    written = [os.path.join(a.out, "info_passing_time.svg")]
    info_passing_figure(written[0])
    if a.results:
        with open(a.results) as fh:
            written += sweep_figures(json.load(fh), a.out)
    if a.curves:
        from .report import accuracy_curves
        with open(a.curves) as fh:
            p = os.path.join(a.out, "accuracy_vs_round.svg")
            line_chart_svg(accuracy_curves(json.load(fh)), "Global accuracy vs round", "accuracy", p)
            written.append(p)
    for p in written:
        print(p)
    return written


if __name__ == "__main__":
    main()
