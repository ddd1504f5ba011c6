#!/bin/bash
# round 2, first GPU pass: the whole GPU suite (new full-width / LM-head / tiling / flash-attn / feeder cases), then the bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r2a_gpu.txt
python -m pytest tests -m gpu -q -s --durations=15 > gpurun_out/r2a_tests.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2a_tests.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
echo "bench exit $?" >> gpurun_out/r2a_bench.err
tail -5 gpurun_out/r2a_tests.log; tail -c 1500 gpurun_out/r2a_bench.json
