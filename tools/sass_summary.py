#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-native SASS mnemonics in libplaincv_b200.so (cuobjdump -sass): UTCHMMA (tcgen05.mma,
`.2CTA` = cta_group::2), LDTM / STTM (tcgen05.ld / st), UTMALDG / UTMASTG / UTMAREDG (TMA load / store / reduce-add),
plus the legacy HMMA.16816 that must not appear.  Written to profiles/rNN_sass_summary.txt by __graft_entry__.build(),
so the evidence is committed even though the .so is not.
usage: python tools/sass_summary.py [lib.so] [out.txt]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MNEMONICS = ("UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTCBAR", "MUFU.EX2",
             "HMMA.16816")


def demangle(names):
    try:
        out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout
        return out.splitlines()
    except Exception:
        return list(names)


def summarize(lib):
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    counts = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        counts[cur]["_instr"] += 1
        for mn in MNEMONICS:
            if op == mn or op.startswith(mn + "."):
                counts[cur][mn] += 1
    return counts


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "plaincv_b200", "lib", "libplaincv_b200.so")
    out = sys.argv[2] if len(sys.argv) > 2 else None
    counts = summarize(lib)
    names = demangle(list(counts))
    cols = [m for m in MNEMONICS]
    lines = [f"# cuobjdump -sass {os.path.relpath(lib, ROOT)} (sm_100a): instruction counts per kernel",
             f"# {'kernel':88s} " + " ".join(f"{c:>12s}" for c in ["instr"] + cols)]
    tot = collections.Counter()
    for (mangled, c), name in zip(counts.items(), names):
        short = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "")).replace("void ", "")
        lines.append(f"{short[:90]:90s} " + " ".join(f"{c[k]:12d}" for k in ["_instr"] + cols))
        tot.update(c)
    lines.append(f"{'TOTAL':90s} " + " ".join(f"{tot[k]:12d}" for k in ["_instr"] + cols))
    text = "\n".join(lines) + "\n"
    if out:
        with open(out, "w") as f:
            f.write(text)
    else:
        sys.stdout.write(text)
    return tot


if __name__ == "__main__":
    main()
