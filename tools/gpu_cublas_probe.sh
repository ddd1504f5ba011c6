#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 300 python tools/bench_gemm_graph.py > $O/r02_gemm_graph.txt 2>&1; echo "gemm graph exit $?"; cat $O/r02_gemm_graph.txt | tail -14
timeout 120 python tools/ncu_cublas_probe.py > $O/cublas_probe_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none -k regex:'nvjet|cutlass|xmma|gemm|sm100|sm90|cublas' -f -o $O/cublas_probe \
    python tools/ncu_cublas_probe.py > $O/cublas_probe_ncu.log 2>&1; echo "ncu exit $?"
ncu -i $O/cublas_probe.ncu-rep --page raw --csv > $O/cublas_probe_raw.csv 2> /dev/null; rm -f $O/cublas_probe.ncu-rep
tail -3 $O/cublas_probe_ncu.log
