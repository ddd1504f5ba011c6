#!/bin/bash
# 2-GPU bench: layers per gradient bucket (fewer, larger all-reduces; tapered groups) and the wire dtype
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline 2> gpurun_out/n2_$tag.err | tee gpurun_out/n2_$tag.json | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag', d['value'], d['ms_per_step'], d['e2e']['value'])"; }
run g1 PCV_DDP_LAYERS_PER_BUCKET=1
run g8taper PCV_DDP_LAYERS_PER_BUCKET=8
run g4taper PCV_DDP_LAYERS_PER_BUCKET=4
