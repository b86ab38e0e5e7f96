"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections, csv, re, sys
path = sys.argv[1]
lines = [l for l in open(path) if not l.startswith("==")]
agg, tot = collections.defaultdict(lambda: [0, 0.0]), 0.0
for row in csv.DictReader(lines):
    v = float(row["Metric Value"].replace(",", ""))
    v = v / 1e3 if row["Metric Unit"] == "ns" else (v * 1e3 if row["Metric Unit"] == "ms" else v)
    name = re.sub(r"\(.*", "", row["Kernel Name"])
    name = re.sub(r"void |fedb::|at::native::", "", name)[:64]
    agg[name][0] += 1
    agg[name][1] += v
    tot += v
print(f"total {tot:.1f} us over {sum(a[0] for a in agg.values())} launches")
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:22]:
    print(f"{t:10.1f} us {100 * t / tot:5.1f}%  n={n:4d} avg {t / n:8.1f}  {k}")
