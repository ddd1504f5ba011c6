#!/bin/bash
# attention A/B: library variants under tools/ab/, CUDA-event timings at the C3 shape (L2 flushed between launches)
mkdir -p gpurun_out
python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "attention" > gpurun_out/attn_tests.log 2>&1; echo "attention tests exit $?"; tail -2 gpurun_out/attn_tests.log
for lib in tools/ab/*.so; do
  echo "== $lib"
  PCV_KB_LIB=$lib python tools/bench_kernels.py 2048 attn 2>&1 | grep "^attn"
done | tee gpurun_out/attn_ab.txt
if [ -f tools/libpcv_trace.so ]; then python tools/trace_attn.py > gpurun_out/attn_trace.txt 2>&1; head -45 gpurun_out/attn_trace.txt; fi
