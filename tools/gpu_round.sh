#!/bin/bash
# one call: the whole GPU suite, the bench line (all configs), and the per-kernel microbench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench exit $?"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_c3.json"))
    print("C3", round(d["value"]), "tok/s", round(d["ms_per_step"], 2), "ms  mfu", round(d["mfu"]["of_burst_peak"], 4), "e2e", round(d["e2e"]["value"]), d["clocks"])
    print("in_step", json.dumps(d["roofline"]["in_step"].get("by_kernel")), "idle", d["roofline"]["in_step"].get("idle_ms"))
    for k, v in d.get("configs", {}).items():
        print(k, round(v["value"]), round(v["ms_per_step"], 2), round(v["mfu"]["of_burst_peak"], 4))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/bench_c3.err").read()[-2000:])
PY
