#!/bin/bash
# Round 2, GPU session H (1 GPU): full parity suite on the retuned classes, alternating-sweep shape sweep at 2^28,
# the per-size table (b200 vs the reference's cuda and thrust models, all through the unmodified driver), bench.py.
set -u
O=gpurun_out/r2h
mkdir -p $O
drv() { "$1" --csv --gigabytes -s $2 --only $3 -n $4 "${@:5}" 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | awk -F, '{printf "%s %.2f  ", $1, $2*1e6; s+=$2*1e6} END {printf "sum %.2f", s}'; }
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_gpu.log
timeout 600 python tools/sweep.py --kernels mul,copy,add,triad --log2n 28 --blocking --alternate --hints 1,2,3,4 --top 16 --out $O/sweep_alt_2p28.csv > $O/sweep_alt_2p28.log 2>&1; grep -A9 "^==" $O/sweep_alt_2p28.log
echo "== per-size table: min us per kernel through src/main.cpp" | tee $O/size_table.txt
for s in 18 20 21 22 23 24 25 26 27 28 29 30; do
  n=$((1 << s)); it=100; [ $s -ge 27 ] && it=40; [ $s -ge 29 ] && it=20
  echo "f64 2^$s b200:   $(drv build/b200-stream $n All $it)" | tee -a $O/size_table.txt
  echo "f64 2^$s thrust: $(drv oracle/_ref/thrust-stream $n All $it)" | tee -a $O/size_table.txt
  echo "f64 2^$s cuda:   $(drv oracle/_ref/cuda-stream $n All $it)" | tee -a $O/size_table.txt
done
for s in 20 24 26 28 30; do
  n=$((1 << s)); it=100; [ $s -ge 28 ] && it=40
  echo "f32 2^$s b200:   $(drv build/b200-stream $n All $it --float)" | tee -a $O/size_table.txt
  echo "f32 2^$s thrust: $(drv oracle/_ref/thrust-stream $n All $it --float)" | tee -a $O/size_table.txt
  echo "f32 2^$s cuda:   $(drv oracle/_ref/cuda-stream $n All $it --float)" | tee -a $O/size_table.txt
done
n=$((1 << 29))
for i in 1 2 3; do
  echo "2^29 Triad-only b200: $(drv build/b200-stream $n Triad 100)   thrust: $(drv oracle/_ref/thrust-stream $n Triad 100)   cuda: $(drv oracle/_ref/cuda-stream $n Triad 100)" | tee -a $O/size_table.txt
done
echo "== bench"
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"; cat $O/bench_n1.json; tail -3 $O/bench_n1.err
python bench.py --impl reference > $O/bench_reference.json 2> $O/bench_reference.err; echo "bench ref rc=$?"; cut -c1-600 $O/bench_reference.json
echo "gpu_r2_h done"
