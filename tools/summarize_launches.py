#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
usage: python tools/summarize_launches.py gpurun_out/launches.csv [first_id last_id]
       python tools/summarize_launches.py gpurun_out/launches.csv --step K
--step K keeps exactly the K-th (0-based) training step of the capture: from the gradient zero-fill / bf16 weight
refresh in front of the K-th embed_fwd launch up to that step's embed_bwd (the last kernel of backward), so that the
per-kernel shares are those of one bench.py step and exclude the bench's separate GEMM replay."""
import csv
import re
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    step = None
    if len(sys.argv) > 3 and sys.argv[2] == "--step":
        step = int(sys.argv[3])
        lo, hi = 0, 1 << 60
    else:
        lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
        hi = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 60
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        i = int(r["ID"])
        if i < lo or i > hi:
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ns = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1)
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        name = re.sub(r"<.*", "", name)
        rows.append((i, name, ns))
    if step is not None:
        fwd = [k for k, r in enumerate(rows) if "embed_fwd" in r[1]]
        k0 = fwd[step]
        while k0 > 0 and ("cast_pad" in rows[k0 - 1][1] or "at::" in rows[k0 - 1][1]):
            k0 -= 1
        k1 = next(k for k in range(fwd[step], len(rows)) if "embed_bwd" in rows[k][1])
        rows = rows[k0:k1 + 1]
    agg = defaultdict(lambda: [0, 0.0])
    for _, n, ns in rows:
        agg[n][0] += 1
        agg[n][1] += ns
    tot = sum(v[1] for v in agg.values())
    print(f"launches {len(rows)}  total {tot/1e6:.3f} ms  (ids {rows[0][0]}..{rows[-1][0]})")
    print(f"{'kernel':60s} {'n':>5s} {'ms':>9s} {'share':>7s} {'avg us':>9s}")
    for n, (c, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n[:60]:60s} {c:5d} {ns/1e6:9.3f} {100*ns/tot:6.1f}% {ns/c/1e3:9.1f}")


if __name__ == "__main__":
    main()
