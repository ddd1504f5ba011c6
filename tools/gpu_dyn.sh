#!/bin/bash
# dynamic (cluster-launch-control) tile schedule of the CTA-pair GEMM: parity first (hang-debug build: a stuck barrier
# traps instead of hanging), then the whole suite, then A/B timings against the static schedule
mkdir -p gpurun_out; O=gpurun_out
PCV_GEMM_DYNAMIC=1 PCV_LIB_PATH=tools/ab/libpcv_hang.so timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x > $O/dyn_kernels_hang.log 2>&1
rc=$?; echo "kernels (hang-debug build, dynamic) exit $rc"; tail -3 $O/dyn_kernels_hang.log
if [ $rc -ne 0 ]; then grep -m5 -i "hang\|error\|trap" $O/dyn_kernels_hang.log; tail -30 $O/dyn_kernels_hang.log; exit 0; fi
PCV_GEMM_DYNAMIC=1 timeout 600 python -m pytest tests -m gpu -q -x > $O/dyn_pytest.log 2>&1
rc=$?; echo "pytest (dynamic) exit $rc"; tail -3 $O/dyn_pytest.log
if [ $rc -ne 0 ]; then tail -40 $O/dyn_pytest.log; exit 0; fi
for m in 0 1; do
  PCV_GEMM_DYNAMIC=$m PCV_GG_LM=1 timeout 300 python tools/bench_gemm_graph.py > $O/dyn_gg$m.txt 2>&1; echo "gemm graph dynamic=$m exit $?"
done
paste -d'\n' $O/dyn_gg0.txt $O/dyn_gg1.txt | cut -c1-60
for m in 0 1 0 1; do
  PCV_GEMM_DYNAMIC=$m timeout 300 python bench.py --steps 20 --warmup 3 --no-extra-configs --no-cpu-baseline --no-in-step-profile > $O/dyn_bench$m.json 2> $O/dyn_bench$m.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("$O/dyn_bench$m.json") if l.startswith("{")][-1])
    print("dynamic=$m", round(d["value"]), round(d["ms_per_step"], 3), round(d["e2e"]["value"]), d["clocks"]["sm_mhz"], round(d["roofline"]["gemm_ms_per_step"], 2), round(d["accum4"]["value"]))
except Exception as e:
    print("failed", e); print(open("$O/dyn_bench$m.err").read()[-2000:])
PY
done
