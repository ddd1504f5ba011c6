#!/bin/bash
# One gpurun call: GPU parity tests, the bench line of every BASELINE config, the launch list of the bench command.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest=$?"
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench=$?"
for c in c2 c4 c5; do
  timeout 300 python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "bench_$c=$?"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu1=$?"
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_c3.json; cat gpurun_out/bench_c[245].json
