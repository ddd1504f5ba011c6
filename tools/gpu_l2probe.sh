#!/bin/bash
# Is the CTA-pair GEMM mainloop bound by the L2 -> SM operand feed?  Same launches with the B / A / both operand loads left out.
for dbg in 0 64 128 192; do
  echo "== PCV_GEMM_DBG=$dbg"
  PCV_GEMM_DBG=$dbg python tools/bench_kernels.py 2048 gemm 2>&1 | grep "^gemm_" 
done
