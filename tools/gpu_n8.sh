#!/bin/bash
# 8-GPU A/B of the gradient all-reduce: overlapped vs after backward, NCCL CTA limits
mkdir -p gpurun_out
run() {
  tag=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \
      bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/n8_$tag.json 2> gpurun_out/n8_$tag.err
  python - <<PY
import json
for l in open("gpurun_out/n8_$tag.json"):
    if l.startswith("{"):
        d = json.loads(l); print("$tag", round(d["value"]), round(d["ms_per_step"], 2), d["clocks"]["sm_mhz"])
PY
}
run base X=1
run nooverlap PCV_DDP_OVERLAP=0
run cta8 NCCL_MAX_CTAS=8
run cta4 NCCL_MAX_CTAS=4
run info NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL
grep -i -m5 "nvls" gpurun_out/n8_info.err gpurun_out/n8_info.json | cut -c1-200
grep -i -m3 "channels\|Algo" gpurun_out/n8_info.err | cut -c1-200
