#!/bin/bash
# final check of the round at HEAD: smoke, the GPU suite, the default bench line, the reference arm, the GGN matvec bench
mkdir -p gpurun_out; O=gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/v_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/v_smoke.log
timeout 600 python -m pytest tests -m gpu -q -x > $O/v_pytest.log 2>&1; echo "pytest exit $?"; tail -2 $O/v_pytest.log
timeout 600 python bench.py > $O/v_bench.json 2> $O/v_bench.err; echo "bench exit $?"
python - <<PY
import json
try:
    d = json.loads([l for l in open("$O/v_bench.json") if l.startswith("{")][-1])
    print(round(d["value"]), round(d["ms_per_step"], 2), round(d["e2e"]["value"]), d["gpu_launches"], d["clocks"], d["roofline"]["frac"], d["mfu"])
    print({k: (round(v["value"]), v["clocks"]["samples"]) for k, v in d["configs"].items()}, d["cpu_baseline"])
except Exception as e:
    print("failed", e); print(open("$O/v_bench.err").read()[-3000:])
PY
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > $O/v_bench_ref.json 2> $O/v_bench_ref.err; echo "reference arm exit $?"; tail -1 $O/v_bench_ref.json | cut -c1-400
timeout 300 python tools/bench_curvature.py > $O/v_curvature.log 2>&1; echo "curvature exit $?"; tail -4 $O/v_curvature.log | cut -c1-400
