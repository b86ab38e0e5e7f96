"""Split an .ncu-rep (ncu --set full) into profiles/ncu/<kernel>.txt: headline metrics of the first launch of every kernel
name + its five hottest SASS lines with their stall reasons.   python bench/ncu_per_kernel.py gpurun_out/prof_all.ncu-rep"""
import csv, io, os, re, subprocess, sys
rep = sys.argv[1]
out_dir = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu")
os.makedirs(out_dir, exist_ok=True)
WANT = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_reads.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_writes.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed_pipe_uniform.sum",
        "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout.split('"Kernel Name",')[1:]
seen = {}
for k, r in enumerate(rows[2:]):
    full = r[idx["Kernel Name"]]
    name = re.sub(r"\(.*", "", full)
    name = re.sub(r"void |fedb::", "", name).strip()
    stem = re.sub(r"[<>, :]+", "_", name).strip("_")
    if stem in seen or name.startswith("at::") or "elementwise_kernel" in name or "cub::" in name or "vectorized" in name:
        continue
    seen[stem] = k
    lines = [f"# {full[:200]}", f"# ncu --set full --clock-control none, first launch of this kernel in bench/ncu_kernels.py (B200, sm_100a)", ""]
    for w in WANT:
        if w in idx:
            lines.append(f"{w:80s} {r[idx[w]]:>16s} {units[idx[w]]}")
    if k < len(src):
        b = src[k].split("\n")
        rd = csv.reader(io.StringIO("\n".join(b[1:]))); h = next(rd)
        iS, iN = h.index("Source"), h.index("# Samples")
        st = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
        rs = [x for x in rd if len(x) > iN]
        tot = sum(int(x[iN] or 0) for x in rs) or 1
        agg = {}
        for x in rs:
            for i in st:
                agg[h[i][6:]] = agg.get(h[i][6:], 0) + int(x[i] or 0)
        lines += ["", "stall samples: " + ", ".join(f"{a}={v}" for a, v in sorted(agg.items(), key=lambda t: -t[1]) if v), "", "hottest SASS lines:"]
        for j, x in sorted(enumerate(rs), key=lambda t: -int(t[1][iN] or 0))[:6]:
            lines.append(f"  {100 * int(x[iN] or 0) / tot:5.1f}%  {x[iS].strip()[:120]}")
    open(os.path.join(out_dir, stem[:80] + ".txt"), "w").write("\n".join(lines) + "\n")
print(f"wrote {len(seen)} kernel summaries to {out_dir}")
