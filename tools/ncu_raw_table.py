#!/usr/bin/env python
"""One line per captured launch from `ncu -i rep --page raw --csv`: duration, clocks, pipe utilisation, DRAM bytes.
usage: python tools/ncu_raw_table.py gpurun_out/kernels_full_raw.csv > profiles/rNN_ncu_full_kernels.txt"""
import csv
import sys

WANT = [("gpu__time_duration.sum", "dur us"), ("sm__cycles_elapsed.avg.per_second", "sm GHz"),
        ("sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed", "tensor %"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("dram__bytes_read.sum", "dram rd MB"), ("dram__bytes_write.sum", "dram wr MB"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %"), ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid")]
SCALE = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "hz": 1e-9, "Khz": 1e-6, "Mhz": 1e-3, "Ghz": 1.0,
         "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "Tbyte": 1e6}

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
print("# ncu --set full --clock-control none --import-source on, PCV_KB_ITERS=1 PCV_KB_WARM=1 python tools/bench_kernels.py")
print("# (C3 shapes, 1 x B200); one row per captured launch, columns from `--page raw`")
print(f"{'kernel':40s} " + " ".join(f"{n:>10s}" for _, n in WANT))
for r in rows[2:]:
    short = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("pcv::", "")[:40]
    vals = []
    for m, _ in WANT:
        if m not in ix:
            vals.append("-")
            continue
        try:
            f = float(r[ix[m]].replace(",", "")) * SCALE.get(units[ix[m]], 1.0)
            vals.append(f"{f:.2f}" if "per_second" in m else f"{f:.1f}")
        except ValueError:
            vals.append(r[ix[m]] or "-")
    print(f"{short:40s} " + " ".join(f"{v:>10s}" for v in vals))
