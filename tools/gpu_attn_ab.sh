#!/bin/bash
# attention A/B: library variants under tools/ab/, CUDA-event timings at the C3 shape (L2 flushed between launches)
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_kernels_gpu.py tests/test_curvature_gpu.py -m gpu -q -x -k "attention or attn" > gpurun_out/attn_tests.log 2>&1; echo "attention tests exit $?"; tail -3 gpurun_out/attn_tests.log
for lib in tools/ab/*.so; do
  echo "== $lib"
  PCV_KB_LIB=$lib timeout 60 python tools/bench_kernels.py 2048 attn 2>&1 | grep "^attn"
  PCV_KB_LIB=$lib timeout 60 python tools/bench_kernels.py 4096 attn 2>&1 | grep "^attn" | sed 's/^/T4096 /'
done | tee gpurun_out/attn_ab.txt
if [ -f tools/libpcv_trace.so ]; then timeout 60 python tools/trace_attn.py > gpurun_out/attn_trace.txt 2>&1; grep -A4 "CTA occupancy" gpurun_out/attn_trace.txt | cut -c1-400; fi
