#!/bin/bash
# Is the CTA-pair GEMM mainloop bound by the L2 -> SM operand feed?  Same launches with the B / A / both operand loads left out.
# The switches exist only in a measurement build of the library; the product build is restored afterwards.
make -C plaincv_b200/csrc clean >/dev/null && make -C plaincv_b200/csrc -j8 all XFLAGS=-DPCV_MEASURE >/dev/null
for dbg in 0 64 128 192; do
  echo "== PCV_GEMM_DBG=$dbg"
  PCV_GEMM_DBG=$dbg python tools/bench_kernels.py 2048 gemm 2>&1 | grep "^gemm_" 
done
make -C plaincv_b200/csrc clean >/dev/null && make -C plaincv_b200/csrc -j8 all >/dev/null
