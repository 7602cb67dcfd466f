#!/bin/bash
# Last short GPU session of a round: thrust baseline, full bench line, parity suite.
set -u
mkdir -p gpurun_out
for s in 33554432 536870912; do
  timeout 60 ./oracle/_ref/thrust-stream --csv --gigabytes -s $s --only All -n 30 > gpurun_out/drv_ref_thrust_$s.csv 2> gpurun_out/drv_ref_thrust_$s.err; echo "thrust $s rc=$?"
  grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," gpurun_out/drv_ref_thrust_$s.csv | cut -d, -f1,5 | tr '\n' ' '; echo
done
timeout 200 python bench.py > gpurun_out/bench_last.json 2> gpurun_out/bench_last.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_last.json
timeout 150 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_last.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_last.log
