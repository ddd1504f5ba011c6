#!/bin/bash
# One gpurun call: GPU parity tests, bench, per-kernel microbench, ncu launch list + full capture of every pcv kernel.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest=$?"
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench=$?"
python tools/bench_kernels.py > gpurun_out/kbench.log 2>&1; echo "kbench=$?"
if [ "$1" != "noncu" ]; then
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu1=$?"
PCV_KB_ITERS=1 PCV_KB_WARM=1 python tools/bench_kernels.py > gpurun_out/plain2.log 2>&1 && \
PCV_KB_ITERS=1 PCV_KB_WARM=1 ncu --set full --clock-control none --import-source on \
    -k regex:'gemm_bf16|attn_|rmsnorm|swiglu|softmax_ce' -f -o gpurun_out/kernels_full \
    python tools/bench_kernels.py > gpurun_out/ncu_full.log 2>&1; echo "ncu2=$?"
# the report itself (source import, ~40 replays x 30 kernels) is larger than what gpurun copies back: keep its CSV pages
ncu -i gpurun_out/kernels_full.ncu-rep --page raw --csv > gpurun_out/kernels_full_raw.csv 2> /dev/null
ncu -i gpurun_out/kernels_full.ncu-rep --page details --csv > gpurun_out/kernels_full_details.csv 2> /dev/null
ls -la gpurun_out/kernels_full.ncu-rep; rm -f gpurun_out/kernels_full.ncu-rep
# DRAM traffic of one step's GEMM launches (roofline.traffic): every GEMM of a block and of the LM head once
PCV_KB_WARM=1 python tools/bench_kernels.py 2048 gemmstep > gpurun_out/plain3.log 2>&1 && \
PCV_KB_WARM=1 ncu --set full --clock-control none -k regex:'gemm_bf16' -f -o gpurun_out/gemmstep \
    python tools/bench_kernels.py 2048 gemmstep > gpurun_out/ncu_gemmstep.log 2>&1; echo "ncu3=$?"
ncu -i gpurun_out/gemmstep.ncu-rep --page raw --csv > gpurun_out/gemmstep_raw.csv 2> /dev/null
rm -f gpurun_out/gemmstep.ncu-rep
fi
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_c3.json
