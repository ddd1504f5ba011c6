#!/bin/bash
# round 2: CUDA-graph step A/B on one GPU + the new tests
mkdir -p gpurun_out
python -m pytest tests/test_model_gpu.py -m gpu -q -x -k "cuda_graph or feeder" > gpurun_out/r2b_tests.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2b_tests.log
PCV_CUDA_GRAPH=0 python bench.py --steps 10 --warmup 3 --no-extra-configs --no-cpu-baseline > gpurun_out/r2b_bench_eager.json 2> gpurun_out/r2b_bench_eager.err
python bench.py --steps 10 --warmup 3 --no-extra-configs --no-cpu-baseline > gpurun_out/r2b_bench_graph.json 2> gpurun_out/r2b_bench_graph.err
tail -3 gpurun_out/r2b_tests.log
for f in eager graph; do python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2b_bench_$f.json"))
    print("$f", d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["in_step"].get("idle_ms"), d["roofline"]["in_step"].get("span_ms"))
except Exception as e:
    print("$f failed", e); print(open("gpurun_out/r2b_bench_$f.err").read()[-2000:])
PY
done
