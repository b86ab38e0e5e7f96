"""Write docs/sass/<kernel>.sass — one SASS listing per kernel of libfedb200.so (``cuobjdump -sass``, sm_100a).

BASELINE.json asks that every kernel be shown in a committed SASS listing.  Tensor-core / TMA kernels are written in full
(that is where ``UTCHMMA`` = tcgen05.mma, ``LDTM`` = tcgen05.ld, ``UTMALDG``/``UTMASTG`` = TMA, ``UTCBAR`` = tcgen05.commit,
``SYNCS`` = mbarrier have to be seen in context); streaming kernels get a header with mnemonic counts plus every line that
touches global memory, atomics, fences or MUFU, so peer loads/stores (``LDG``/``STG`` on mapped pointers), system-scope
fences (``MEMBAR.SC.SYS``, ``LD.E.STRONG.SYS``) and reductions are visible without thousands of ALU lines.  Template
instances of one kernel share a file name stem and are numbered.  Regenerate after a rebuild:

    python bench/sass_listings.py            # also rewrites docs/sass/INDEX.md
"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fedb200", "lib", "libfedb200.so")
OUT = os.path.join(ROOT, "docs", "sass")
KEY = re.compile(r"\b(UTCHMMA|UTCQMMA|UTCBAR|UTCCP|LDTM|STTM|UTMALDG|UTMASTG|UTMAPF|UBLKCP|SYNCS|HMMA|REDG|RED|ATOMG|ATOM|MUFU|"
                 r"LDGMC|STGMC|REDGMC|LDG|STG|LD|ST|MEMBAR|ERRBAR|CCTL|NANOSLEEP|ACQBULK|UCGABAR|BAR)\b")
FULL_IF = ("UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "UBLKCP")
MAX_FULL_LINES = 12000


def demangle(n: str) -> str:
    r = subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
    return r.replace("fedb::", "")


def main() -> None:
    os.makedirs(OUT, exist_ok=True)
    for f in os.listdir(OUT):
        if f.endswith(".sass") or f == "INDEX.md":
            os.remove(os.path.join(OUT, f))
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    funcs, cur = collections.OrderedDict(), None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur is not None and re.match(r"\s+/\*[0-9a-f]{4,6}\*/", line):
            funcs[cur].append(re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", line.rstrip()))
    seen, index = collections.Counter(), []
    for mangled, lines in funcs.items():
        full = demangle(mangled)
        stem = re.sub(r"[<(].*", "", full).replace("void ", "").strip()
        seen[stem] += 1
        name = f"{stem}.sass" if seen[stem] == 1 else f"{stem}.{seen[stem]}.sass"
        counts = collections.Counter(m.group(1) for l in lines for m in [KEY.search(l)] if m)
        tensor = any(counts[k] for k in FULL_IF)
        # full text for the first instance of every tensor-core kernel; further template instances differ only in their
        # epilogue and get the excerpt form
        body = lines if (tensor and seen[stem] == 1 and len(lines) <= MAX_FULL_LINES) else [l for l in lines if KEY.search(l)]
        kind = "full listing" if body is lines else "excerpt: memory / tensor / sync / MUFU instructions only"
        with open(os.path.join(OUT, name), "w") as fh:
            fh.write(f"// {full}\n// {len(lines)} SASS instructions ({kind}); sm_100a, cuobjdump -sass\n")
            fh.write("// " + ", ".join(f"{k}={v}" for k, v in sorted(counts.items()) if v) + "\n")
            fh.write("\n".join(body) + "\n")
        index.append((name, full, len(lines), counts))
    with open(os.path.join(OUT, "INDEX.md"), "w") as fh:
        fh.write("# SASS listings (sm_100a) — regenerate with `python bench/sass_listings.py`\n\n"
                 "`UTCHMMA(.2CTA)` = tcgen05.mma, `LDTM` = tcgen05.ld, `UTMALDG`/`UTMASTG` = TMA load/store, `UTCBAR` = "
                 "tcgen05.commit, `SYNCS` = mbarrier; no `HMMA` (mma.sync) anywhere.\n\n"
                 "| file | kernel | instr | UTCHMMA | LDTM | UTMALDG | UTMASTG | UBLKCP | SYNCS | RED/ATOM | MUFU | HMMA |\n"
                 "|---|---|---|---|---|---|---|---|---|---|---|---|\n")
        for name, full, n, c in index:
            short = re.sub(r"\(.*", "", full)
            fh.write(f"| [{name}]({name}) | `{short[:90]}` | {n} | {c['UTCHMMA']} | {c['LDTM']} | {c['UTMALDG']} | {c['UTMASTG']} | "
                     f"{c['UBLKCP']} | {c['SYNCS']} | {c['REDG'] + c['RED'] + c['ATOM'] + c['ATOMG']} | {c['MUFU']} | {c['HMMA']} |\n")
    print(f"wrote {len(index)} listings to docs/sass/")


if __name__ == "__main__":
    main()
