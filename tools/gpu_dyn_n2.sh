#!/bin/bash
# 2 GPUs: static vs dynamic GEMM tile schedule under the gradient all-reduce, then the torchrun gradient-equality tests
N=${1:-2}
mkdir -p gpurun_out; O=gpurun_out
for m in 0 1 0 1; do
  PCV_GEMM_DYNAMIC=$m timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 20 --warmup 3 --no-extra-configs --no-cpu-baseline --no-in-step-profile --no-ddp-check > $O/dyn_n${N}_bench$m.json 2> $O/dyn_n${N}_bench$m.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("$O/dyn_n${N}_bench$m.json") if l.startswith("{")][-1])
    print("N=$N dynamic=$m", round(d["value"]), round(d["ms_per_step"], 3), round(d["e2e"]["value"]), d["clocks"]["sm_mhz"], round(d["accum4"]["value"]))
except Exception as e:
    print("failed", e); print(open("$O/dyn_n${N}_bench$m.err").read()[-2000:])
PY
done
PCV_GEMM_DYNAMIC=1 timeout 900 python -m pytest tests/test_ddp_gpu.py -m gpu -q -s > $O/dyn_n${N}_ddp_tests.log 2>&1
echo "ddp pytest (dynamic) exit $?"; tail -2 $O/dyn_n${N}_ddp_tests.log
