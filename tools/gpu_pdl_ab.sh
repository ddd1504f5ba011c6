#!/bin/bash
# programmatic dependent launch on / off, same box: tests once, then the C3 bench line twice each (alternating)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
for rep in 1 2; do
for pdl in 0 1; do
  PCV_PDL=$pdl timeout 300 python bench.py --steps 20 --warmup 3 --no-extra-configs --no-cpu-baseline > gpurun_out/bench_pdl$pdl.json 2> gpurun_out/bench_pdl$pdl.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_pdl$pdl.json"))
    print("PDL=$pdl", round(d["value"]), "tok/s", round(d["ms_per_step"], 3), "ms e2e", round(d["e2e"]["value"]), "idle", round(d["roofline"]["in_step"].get("idle_ms", -1), 3), "span", round(d["roofline"]["in_step"].get("span_ms", -1), 3), d["clocks"]["sm_mhz"])
except Exception as e:
    print("PDL=$pdl failed", e); print(open("gpurun_out/bench_pdl$pdl.err").read()[-1500:])
PY
done
done
