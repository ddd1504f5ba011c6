#!/bin/bash
# round 2, N GPUs: the bench as the driver launches it (ddp_check inside), without the extra configs / CPU baseline
N=${1:-8}
mkdir -p gpurun_out; O=gpurun_out
timeout 700 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 20 --warmup 3 --no-extra-configs --no-cpu-baseline > $O/r02_n${N}_bench.json 2> $O/r02_n${N}_bench.err
echo "bench N=$N exit $?"
python - <<PY
import json
try:
    d = json.loads([l for l in open("$O/r02_n${N}_bench.json") if l.startswith("{")][-1])
    print("N=$N", round(d["value"]), round(d["ms_per_step"], 2), round(d["e2e"]["value"]), d["clocks"]["sm_mhz"], round(d["accum4"]["value"]), json.dumps(d.get("ddp_check"))[:900])
except Exception as e:
    print("failed", e); print(open("$O/r02_n${N}_bench.err").read()[-3000:])
PY
