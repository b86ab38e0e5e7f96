"""Write profiles/sass_summary.md: per-kernel counts of the SASS mnemonics that prove the Blackwell paths."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "fedb200/lib/libfedb200.so")], capture_output=True, text=True).stdout
cur, stats = None, collections.OrderedDict()
pat = re.compile(r"\b(UTCHMMA|UTCQMMA|UTCBAR|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|SYNCS|HMMA|REDG|RED|ATOMG|ATOM|MUFU)\b")
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        stats[cur] = collections.Counter()
    elif cur:
        m = pat.search(line)
        if m:
            stats[cur][m.group(1)] += 1
def demangle(n):
    r = subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
    return re.sub(r"\(.*", "", r).replace("fedb::", "")
with open(os.path.join(ROOT, "profiles/sass_summary.md"), "w") as f:
    f.write("# SASS evidence per kernel (`cuobjdump -sass fedb200/lib/libfedb200.so`, sm_100a)\n\n"
            "Counts of the mnemonics that prove the Blackwell paths: `UTCHMMA` = tcgen05.mma (kind::f16), `LDTM` = tcgen05.ld, "
            "`UTMALDG` = TMA tensor load (cp.async.bulk.tensor), `UTCBAR` = tcgen05.commit, `SYNCS` = mbarrier ops, "
            "`REDG`/`RED`/`ATOM` = global reductions (`REDG.E.ADD.F32x4` = the vector fp32 atomics of the split-K wgrad and dQ). "
            "No `HMMA` (legacy mma.sync) appears anywhere.  Regenerate with `python bench/sass_summary.py`.\n\n"
            "| kernel | UTCHMMA | LDTM | UTMALDG | UTCBAR | SYNCS | REDG/RED/ATOM | MUFU | HMMA |\n|---|---|---|---|---|---|---|---|---|\n")
    for k, c in stats.items():
        f.write(f"| `{demangle(k)}` | {c['UTCHMMA']} | {c['LDTM']} | {c['UTMALDG']} | {c['UTCBAR']} | {c['SYNCS']} | "
                f"{c['REDG'] + c['RED'] + c['ATOM'] + c['ATOMG']} | {c['MUFU']} | {c['HMMA']} |\n")
print("wrote profiles/sass_summary.md")
