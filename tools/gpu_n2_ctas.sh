#!/bin/bash
# 2-GPU bench, default NCCL CTA count against capped ones (does NCCL's SM use slow the persistent GEMMs?)
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline 2> gpurun_out/n2_$tag.err | tee gpurun_out/n2_$tag.json | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag', d['value'], d['ms_per_step'], d['e2e']['value'])"; }
run default FOO=1
run ctas4 NCCL_MAX_CTAS=4
run ctas2 NCCL_MAX_CTAS=2 NCCL_MIN_CTAS=1
