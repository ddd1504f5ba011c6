#!/bin/bash
# correctness of the halved tail tiles (standalone GEMM check + GPU suite), then A/B of the tail in kernels and bench
mkdir -p gpurun_out
timeout 300 ./plaincv_b200/csrc/test_gemm > gpurun_out/test_gemm.log 2>&1; echo "test_gemm=$?"; tail -32 gpurun_out/test_gemm.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest=$?"; tail -3 gpurun_out/pytest_gpu.log
PCV_KB_CUBLAS=0 python tools/bench_kernels.py 2048 gemm > gpurun_out/kbench_tail1.log 2>&1; echo "kb1=$?"
PCV_GEMM_TAIL=0 PCV_KB_CUBLAS=0 python tools/bench_kernels.py 2048 gemm > gpurun_out/kbench_tail0.log 2>&1; echo "kb0=$?"
paste gpurun_out/kbench_tail1.log gpurun_out/kbench_tail0.log | cut -c1-60,85-130
for i in 1 0 1 0; do
PCV_GEMM_TAIL=$i python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tail$i.json 2>&1
python -c "
import json
for l in open('gpurun_out/bench_tail$i.json'):
    if l.startswith('{'):
        d=json.loads(l); print('tail$i', d['value'], d['ms_per_step'], d['roofline']['gemm_ms_per_step'], d['clocks'])"
done
