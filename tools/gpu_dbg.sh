#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "swiglu" 2>&1 | tail -2
PCV_KB_CUBLAS=0 timeout 120 python tools/bench_kernels.py 2048 fused 2>&1 | grep "#1"
timeout 300 python tools/prof_epilogue.py 2>&1 | tail -4
