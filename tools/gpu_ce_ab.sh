#!/bin/bash
# LM head + cross-entropy: un-fused (PCV_HEAD_CE_FUSED=0) against fused, C3 bench line, alternating on one box
mkdir -p gpurun_out
for rep in 1 2; do
for fused in 0 1; do
  PCV_HEAD_CE_FUSED=$fused timeout 300 python bench.py --steps 20 --warmup 3 --no-extra-configs --no-cpu-baseline > gpurun_out/bench_ce$fused.json 2> gpurun_out/bench_ce$fused.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_ce$fused.json"))
    bk = d["roofline"]["in_step"].get("by_kernel", {})
    print("FUSED=$fused", round(d["value"]), "tok/s", round(d["ms_per_step"], 3), "ms loss", bk.get("loss"), "other", bk.get("other"), "gemm", bk.get("gemm"), d["clocks"]["sm_mhz"])
except Exception as e:
    print("FUSED=$fused failed", e); print(open("gpurun_out/bench_ce$fused.err").read()[-1500:])
PY
done
done
