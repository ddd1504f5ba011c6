#!/bin/bash
# round 2, 2 GPUs: the torchrun gradient-equality tests, the bench exactly as the driver launches it (ddp_check inside),
# and the fp32-wire A/B of the gradient all-reduce
N=2
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_ddp_gpu.py -m gpu -q -s > $O/r02_n2_ddp_tests.log 2>&1
echo "pytest exit $?" | tee -a $O/r02_n2_ddp_tests.log
tail -3 $O/r02_n2_ddp_tests.log
run() {
  tag=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 20 --warmup 3 $EXTRA > $O/r02_n2_bench_$tag.json 2> $O/r02_n2_bench_$tag.err
  echo "bench $tag exit $?"
  python - <<PY
import json
try:
    d = json.loads([l for l in open("$O/r02_n2_bench_$tag.json") if l.startswith("{")][-1])
    print("$tag", round(d["value"]), round(d["ms_per_step"], 2), round(d["e2e"]["value"]), d["clocks"]["sm_mhz"], json.dumps(d.get("ddp_check"))[:600])
except Exception as e:
    print("failed", e); print(open("$O/r02_n2_bench_$tag.err").read()[-3000:])
PY
}
EXTRA="" run driver X=1
EXTRA="--no-extra-configs --no-cpu-baseline --no-ddp-check" run fp32wire PCV_DDP_COMM=fp32
EXTRA="--no-extra-configs --no-cpu-baseline --no-ddp-check" run bf16wire X=1
