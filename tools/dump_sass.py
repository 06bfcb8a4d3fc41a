"""Regenerates docs/sass/: for every .cu object, the list of kernels with their register / instruction counts and the
counts of the SASS mnemonics that identify the Blackwell paths (UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG / UTMASTG /
UTMAREDG = TMA, LDGMC / STGMC... = multimem, SYNCS = mbarrier), plus a per-kernel excerpt of the lines that carry them.
Runs on the CPU box (cuobjdump only needs the objects)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "fedml_b200", "native", "obj")
OUT = os.path.join(ROOT, "docs", "sass")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "UTMAPF", "UTCBAR", "SYNCS",
        "LDGMC", "MULTIMEM", "RED.E.ADD", "ATOMG", "ATOM.E", "REDG", "LDG.E.128", "STG.E.128", "LD.E.128.STRONG.SYS", "ST.E.128.STRONG.SYS",
        "LDG.E.STRONG.SYS", "ST.E.STRONG.SYS", "MEMBAR", "ACQBULK", "UCGABAR", "SHFL", "MUFU", "HMMA", "FFMA", "IMAD.WIDE", "ELECT",
        "CCTL", "ERRBAR", "NANOSLEEP", "S2UR", "CS2R"]
os.makedirs(OUT, exist_ok=True)
index = ["# SASS evidence (sm_100a), one file per translation unit — regenerate with `python tools/dump_sass.py`", ""]
for obj in sorted(f for f in os.listdir(OBJ) if f.endswith(".o")):
    r = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, obj)], capture_output=True, text=True)
    if r.returncode != 0:
        continue
    kernels, cur = collections.OrderedDict(), None
    for line in r.stdout.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = []
            continue
        if cur is not None and re.search(r"/\*[0-9a-f]{4}\*/", line):
            kernels[cur].append(line.rstrip())
    dem = subprocess.run(["cu++filt"] + list(kernels), capture_output=True, text=True).stdout.splitlines() if kernels else []
    name = obj[:-2]
    lines = [f"# {name}.cu — cuobjdump -sass, sm_100a: kernels, instruction counts, identifying mnemonics", ""]
    tu_counts = collections.Counter()
    for (k, body), d in zip(kernels.items(), dem or kernels):
        cnt = collections.Counter()
        hits = []
        for ln in body:
            op = ln.split("*/", 1)[1].strip() if "*/" in ln else ln
            for key in KEYS:
                if key in op:
                    cnt[key] += 1
                    if key not in ("FFMA", "IMAD.WIDE", "SHFL", "MUFU", "S2UR", "CS2R") and len(hits) < 40 and \
                            sum(1 for h in hits if key in h) < 3:
                        hits.append(ln.strip())
                    break
        tu_counts.update(cnt)
        lines.append(f"## {d}")
        lines.append(f"{len(body)} SASS instructions; " + ", ".join(f"{k} x{v}" for k, v in cnt.most_common()))
        lines.append("```")
        lines.extend(hits)
        lines.append("```")
        lines.append("")
    with open(os.path.join(OUT, f"{name}.sass.md"), "w") as fh:
        fh.write("\n".join(lines))
    show = [k for k in ("UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "UTMAREDG", "SYNCS", "LDGMC", "LD.E.128.STRONG.SYS", "LDG.E.STRONG.SYS",
                        "ST.E.128.STRONG.SYS", "ERRBAR", "REDG", "ATOMG", "LDG.E.128", "STG.E.128", "SHFL", "FFMA") if tu_counts.get(k)]
    index.append(f"* `{name}.sass.md` — {len(kernels)} kernels; " + ", ".join(f"{k} x{tu_counts[k]}" for k in show))
with open(os.path.join(OUT, "README.md"), "w") as fh:
    fh.write("\n".join(index) + "\n")
print("\n".join(index))
