#!/bin/bash
# round 2, N GPUs: the torchrun gradient-equality tests and the bench (CUDA-graph step, then eager) with its ddp_check
N=${1:-2}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_ddp_gpu.py -m gpu -q -s > gpurun_out/r2n${N}_ddp_tests.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2n${N}_ddp_tests.log
tail -4 gpurun_out/r2n${N}_ddp_tests.log
for mode in 1 0; do
  PCV_CUDA_GRAPH=$mode timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 10 --warmup 3 --no-extra-configs > gpurun_out/r2n${N}_bench_graph${mode}.json 2> gpurun_out/r2n${N}_bench_graph${mode}.err
  echo "bench graph=$mode exit $?"
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/r2n${N}_bench_graph${mode}.json") if l.startswith("{")][-1])
    print("N=$N graph=$mode", d["value"], d["ms_per_step"], d["e2e"]["value"], json.dumps(d.get("ddp_check")))
except Exception as e:
    print("failed", e); print(open("gpurun_out/r2n${N}_bench_graph${mode}.err").read()[-3000:])
PY
done
