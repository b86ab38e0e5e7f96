"""Read an .ncu-rep: per kernel, headline metrics + the SASS instructions that collect the most stall samples."""
import csv, io, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 10
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]; idx = {h: i for i, h in enumerate(hdr)}
want = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size",
        "launch__registers_per_thread", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"]
heads = []
for r in rows[2:]:
    heads.append(r[idx["Kernel Name"]][:60] + " | " + " | ".join(f"{w.split('.')[0][-22:]}={r[idx[w]]}" for w in want if w in idx))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
for bi, b in enumerate(src.split('"Kernel Name",')[1:]):
    lines = b.split("\n")
    print("=" * 100); print(heads[bi] if bi < len(heads) else lines[0][:80])
    rd = csv.reader(io.StringIO("\n".join(lines[1:]))); h = next(rd)
    iS, iN = h.index("Source"), h.index("# Samples")
    rs = [r for r in rd if len(r) > iN]
    tot = sum(int(r[iN] or 0) for r in rs) or 1
    for k, r in sorted(enumerate(rs), key=lambda x: -int(x[1][iN] or 0))[:top]:
        print(f"   #{k:5d} {100 * int(r[iN] or 0) / tot:5.1f}%  {r[iS][:110]}")
    for k, r in enumerate(rs):
        if "TRYWAIT" in r[iS]:
            print(f"   wait #{k:5d} {100 * int(r[iN] or 0) / tot:5.1f}%  {r[iS][:80]}")
