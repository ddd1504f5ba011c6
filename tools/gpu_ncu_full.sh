#!/bin/bash
# One gpurun call: `ncu --set full` of every pcv kernel at the C3 shapes (tools/bench_kernels.py) and of the N4 tangent
# kernels (tools/bench_curvature.py, 2 layers); only the CSV pages travel back.
mkdir -p gpurun_out
PCV_KB_ITERS=1 PCV_KB_WARM=1 timeout 300 python tools/bench_kernels.py > gpurun_out/plain2.log 2>&1 && \
PCV_KB_ITERS=1 PCV_KB_WARM=1 timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:'gemm_bf16|attn_|rmsnorm|swiglu|softmax_ce' -f -o gpurun_out/kernels_full \
    python tools/bench_kernels.py > gpurun_out/ncu_full.log 2>&1; echo "ncu_kernels=$?"
ncu -i gpurun_out/kernels_full.ncu-rep --page raw --csv > gpurun_out/kernels_full_raw.csv 2> /dev/null
rm -f gpurun_out/kernels_full.ncu-rep
timeout 600 ncu --set full --clock-control none -c 60 \
    -k regex:'attn_jvp|rmsnorm_residual_jvp|swiglu_jvp|ce_hessian_vec' -f -o gpurun_out/curv_full \
    python tools/bench_curvature.py 2 > gpurun_out/ncu_curv.log 2>&1; echo "ncu_curv=$?"
ncu -i gpurun_out/curv_full.ncu-rep --page raw --csv > gpurun_out/curv_full_raw.csv 2> /dev/null
rm -f gpurun_out/curv_full.ncu-rep
ls -la gpurun_out/*.csv
