#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "attention" 2>&1 | tail -3
timeout 300 python tools/bench_kernels.py 2048 attn 2>&1 | tail -4
python tools/trace_attn.py 2>&1 | grep -A52 "bwd compute half0 (key"
