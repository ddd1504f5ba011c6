#!/bin/bash
# A/B of the wgrad side stream + ncu capture of the fused-epilogue GEMMs.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest=$?"
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_side1.json 2>&1; echo "bench1=$?"
PCV_WGRAD_STREAM=0 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_side0.json 2>&1; echo "bench0=$?"
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_side1b.json 2>&1; echo "bench1b=$?"
PCV_KB_ITERS=1 PCV_KB_WARM=1 ncu --set full --clock-control none --import-source on -k regex:'pair_kernel' -f -o gpurun_out/fused \
    python tools/bench_kernels.py 2048 fused > gpurun_out/ncu_fused.log 2>&1; echo "ncu=$?"
ncu -i gpurun_out/fused.ncu-rep --page details --csv > gpurun_out/fused_details.csv 2> /dev/null
ncu -i gpurun_out/fused.ncu-rep --page source --csv --kernel-id :::4 > gpurun_out/fused_src_k4.csv 2> /dev/null
ls -la gpurun_out/fused.ncu-rep
tail -3 gpurun_out/pytest_gpu.log
for f in gpurun_out/bench_side1.json gpurun_out/bench_side0.json gpurun_out/bench_side1b.json; do python -c "
import json,sys
for l in open('$f'):
    if l.startswith('{'):
        d=json.loads(l); print('$f', d['value'], d['ms_per_step'], d['clocks'])"; done
