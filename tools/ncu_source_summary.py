#!/usr/bin/env python
"""Summarise `ncu --page source --csv` output (SASS view): opcode histogram by executed instructions and the top
stall-sample instructions.  usage: python tools/ncu_source_summary.py file.csv [topN]"""
import csv
import re
import sys
from collections import defaultdict

path = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path)))
# find header row
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
ops = defaultdict(lambda: [0, 0])
insts = []
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
stall_tot = defaultdict(int)
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or not r[0].startswith("0x"):
        continue
    src = r[col["Source"]].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
    op = m.group(2).split(".")[0] if m else "?"
    ex = int(r[col["Instructions Executed"]] or 0)
    smp = int(r[col["# Samples"]] or 0)
    ops[op][0] += ex
    ops[op][1] += smp
    insts.append((smp, ex, src, r))
    for s in stall_cols:
        stall_tot[s] += int(r[col[s]] or 0)
tot_ex = sum(v[0] for v in ops.values())
tot_s = sum(v[1] for v in ops.values())
print(f"total warp-instructions executed {tot_ex}, samples {tot_s}")
print("opcode histogram (by executed):")
for op, (ex, smp) in sorted(ops.items(), key=lambda kv: -kv[1][0])[:30]:
    print(f"  {op:12s} {ex:12d} {100*ex/tot_ex:5.1f}%   samples {100*smp/max(tot_s,1):5.1f}%")
print("stall reasons (all samples):")
for s, v in sorted(stall_tot.items(), key=lambda kv: -kv[1])[:10]:
    print(f"  {s:28s} {v:8d} {100*v/max(tot_s,1):5.1f}%")
print(f"top {topn} instructions by samples:")
for smp, ex, src, r in sorted(insts, key=lambda t: -t[0])[:topn]:
    top = sorted(((int(r[col[s]] or 0), s) for s in stall_cols), reverse=True)[:2]
    print(f"  {smp:6d} ex={ex:9d}  {src[:70]:70s} {top}")
