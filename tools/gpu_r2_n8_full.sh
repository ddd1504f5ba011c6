#!/bin/bash
# round 2, N GPUs: the bench as the driver launches it, with the C2 / C4 / C5 lines and the N-rank gradient check
N=${1:-8}
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline > $O/r02_n${N}_bench_full.json 2> $O/r02_n${N}_bench_full.err
echo "bench N=$N exit $?"
python - <<PY
import json
try:
    d = json.loads([l for l in open("$O/r02_n${N}_bench_full.json") if l.startswith("{")][-1])
    print("N=$N", round(d["value"]), round(d["ms_per_step"], 2), round(d["e2e"]["value"]), d["clocks"], round(d["accum4"]["value"]), "ddp ok", d["ddp_check"]["ok"])
    for k, v in d["configs"].items():
        print(k, round(v["value"]), round(v["ms_per_step"], 2), v["mfu"], v["clocks"]["samples"])
except Exception as e:
    print("failed", e); print(open("$O/r02_n${N}_bench_full.err").read()[-3000:])
PY
