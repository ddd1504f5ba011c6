#!/bin/bash
PCV_KB_CUBLAS=0 timeout 120 python tools/bench_kernels.py 2048 ew 2>&1 | grep "softmax\|rmsnorm"
PCV_CE_NOCACHE=1 PCV_KB_CUBLAS=0 timeout 120 python tools/bench_kernels.py 2048 ew 2>&1 | grep "softmax"
