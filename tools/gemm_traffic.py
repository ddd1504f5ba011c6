#!/usr/bin/env python
"""DRAM traffic of one training step's GEMM launches from an `ncu --set full` capture of
`python tools/bench_kernels.py 2048 gemmstep` (PCV_KB_WARM=1): the LAST 15 gemm launches of the capture are the 12
GEMMs of one transformer block and the 3 of the LM head.  Writes the JSON bench.py reads for `roofline.traffic`.
usage: python tools/gemm_traffic.py gpurun_out/gemmstep_raw.csv profiles/r01_gemm_traffic.json [n_layers]"""
import csv
import json
import sys

NAMES = ["fwd qkv+rope", "fwd out", "fwd gate|up+swiglu", "fwd down", "dgrad down+dswiglu", "wgrad down", "dgrad gate|up",
         "wgrad gate|up", "dgrad out", "wgrad out", "dgrad qkv", "wgrad qkv", "fwd lm_head", "dgrad lm_head", "wgrad lm_head"]
FLOP = None


def to_bytes(v, unit):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]


def main():
    src, dst = sys.argv[1], sys.argv[2]
    n_layers = int(sys.argv[3]) if len(sys.argv) > 3 else 24
    rows = list(csv.reader(open(src)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    g = [r for r in rows[2:] if "gemm_bf16" in r[ix["Kernel Name"]]][-15:]
    assert len(g) == 15, len(g)
    out = []
    for name, r in zip(NAMES, g):
        rd = to_bytes(r[ix["dram__bytes_read.sum"]], units[ix["dram__bytes_read.sum"]])
        wr = to_bytes(r[ix["dram__bytes_write.sum"]], units[ix["dram__bytes_write.sum"]])
        dur = float(r[ix["gpu__time_duration.sum"]].replace(",", ""))
        du = units[ix["gpu__time_duration.sum"]]
        us = dur * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(du, 1e-3)
        out.append({"gemm": name, "dram_read_bytes": rd, "dram_write_bytes": wr, "duration_us_under_ncu": us})
    layer = sum(o["dram_read_bytes"] + o["dram_write_bytes"] for o in out[:12])
    head = sum(o["dram_read_bytes"] + o["dram_write_bytes"] for o in out[12:])
    res = {"source": "ncu --set full --clock-control none, python tools/bench_kernels.py 2048 gemmstep (C3 shapes, L2 flushed)",
           "n_layers": n_layers, "block_bytes": layer, "head_bytes": head, "step_bytes": n_layers * layer + head,
           "launches": out}
    json.dump(res, open(dst, "w"), indent=1)
    print(json.dumps({k: v for k, v in res.items() if k != "launches"}))
    for o in out:
        print(f"{o['gemm']:22s} read {o['dram_read_bytes']/1e6:9.1f} MB  write {o['dram_write_bytes']/1e6:9.1f} MB  {o['duration_us_under_ncu']:8.1f} us")


if __name__ == "__main__":
    main()
