"""Per-kernel stall-reason totals + hottest SASS lines (with executed counts) from an .ncu-rep."""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 20
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
for b in src.split('"Kernel Name",')[1:]:
    lines = b.split("\n")
    rd = csv.reader(io.StringIO("\n".join(lines[1:]))); h = next(rd)
    iS, iN, iE = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
    rs = [r for r in rd if len(r) > iN]
    tot = sum(int(r[iN] or 0) for r in rs)
    st = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    agg = {}
    for r in rs:
        for i in st: agg[h[i][6:]] = agg.get(h[i][6:], 0) + int(r[i] or 0)
    print("=" * 100); print(lines[0][:90], "samples", tot, "warp-instr", sum(int(r[iE] or 0) for r in rs))
    print({k: v for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v})
    for k, r in sorted(enumerate(rs), key=lambda x: -int(x[1][iN] or 0))[:top]:
        print(f"  #{k:5d} {int(r[iN] or 0):5d} exec {r[iE]:>8s}  {r[iS][:64]:64s}", {h[i][6:]: r[i] for i in st if r[i] not in ("0", "")})
