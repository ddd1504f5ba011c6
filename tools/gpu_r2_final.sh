#!/bin/bash
# round-2 evidence in one call: tests, bench (all configs), per-kernel microbench, ncu launch list, ncu --set full tables,
# GEMM DRAM traffic, compute-sanitizer.  Each ncu / sanitizer command runs only after its plain twin has exited 0.
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > $O/r02_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -2 $O/r02_pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 > $O/r02_bench_c3.json 2> $O/r02_bench_c3.err; echo "bench exit $?"
timeout 300 python tools/bench_kernels.py 2048 all > $O/r02_kernel_microbench_c3.txt 2>&1; echo "microbench exit $?"
timeout 100 python tools/bench_kernels.py 2048 ce >> $O/r02_kernel_microbench_c3.txt 2>&1
# launch list of the bench command
timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra-configs --no-in-step-profile > $O/r02_bench_short.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 5000 --csv --log-file $O/r02_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra-configs --no-in-step-profile > $O/r02_ncu_launches.log 2>&1; echo "ncu launches exit $?"
# ncu --set full, every kernel at the C3 shapes
PCV_KB_ITERS=1 PCV_KB_WARM=1 PCV_KB_CUBLAS=0 timeout 300 python tools/bench_kernels.py 2048 all > /dev/null 2>&1 && \
PCV_KB_ITERS=1 PCV_KB_WARM=1 PCV_KB_CUBLAS=0 timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:'gemm_bf16|attn_|rmsnorm|swiglu|softmax_ce' -f -o $O/kernels_full python tools/bench_kernels.py 2048 all > $O/r02_ncu_full.log 2>&1; echo "ncu full exit $?"
ncu -i $O/kernels_full.ncu-rep --page raw --csv > $O/r02_kernels_full_raw.csv 2> /dev/null; rm -f $O/kernels_full.ncu-rep
PCV_KB_ITERS=1 PCV_KB_WARM=1 timeout 400 ncu --set full --clock-control none -k regex:'gemm_bf16|ce_' -f -o $O/ce_full \
    python tools/bench_kernels.py 2048 ce > $O/r02_ncu_ce.log 2>&1; echo "ncu ce exit $?"
ncu -i $O/ce_full.ncu-rep --page raw --csv > $O/r02_ce_full_raw.csv 2> /dev/null; rm -f $O/ce_full.ncu-rep
# DRAM traffic of one step's GEMM launches
PCV_KB_WARM=1 timeout 600 ncu --set full --clock-control none -k regex:'gemm_bf16' -f -o $O/gemmstep \
    python tools/bench_kernels.py 2048 gemmstep > $O/r02_ncu_gemmstep.log 2>&1; echo "ncu gemmstep exit $?"
ncu -i $O/gemmstep.ncu-rep --page raw --csv > $O/r02_gemmstep_raw.csv 2> /dev/null; rm -f $O/gemmstep.ncu-rep
# compute-sanitizer at small shapes
timeout 120 python tools/sanitize_small.py > $O/r02_sanitize_plain.log 2>&1; echo "sanitize plain exit $?"
for tool in memcheck synccheck racecheck; do
  timeout 420 compute-sanitizer --tool $tool --error-exitcode 7 python tools/sanitize_small.py > $O/r02_sanitizer_$tool.log 2>&1
  echo "sanitizer $tool exit $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize_small: ran" $O/r02_sanitizer_$tool.log | tail -3
done
