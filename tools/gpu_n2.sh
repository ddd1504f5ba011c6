#!/bin/bash
mkdir -p gpurun_out
run() {
  tag=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 \
      bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/n2_$tag.json 2> gpurun_out/n2_$tag.err || tail -5 gpurun_out/n2_$tag.err
  python - <<PY
import json
for l in open("gpurun_out/n2_$tag.json"):
    if l.startswith("{"):
        d = json.loads(l); print("$tag", round(d["value"]), round(d["ms_per_step"], 2), d["clocks"]["sm_mhz"], d["loss"])
PY
}
run bf16 X=1
run fp32 PCV_DDP_COMM=fp32
python bench.py --no-cpu-baseline | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('n1', round(d['value']), round(d['ms_per_step'],2), d['loss'])"
