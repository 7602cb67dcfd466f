#!/bin/bash
# Round 2, GPU session I (1 GPU): dot grid multiples of the resident block count, the fp32 rows of the per-size table, both bench arms.
set -u
O=gpurun_out/r2i
mkdir -p $O
drv() { "$1" --csv --gigabytes -s $2 --only $3 -n $4 "${@:5}" 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | awk -F, '{printf "%s %.2f  ", $1, $2*1e6; s+=$2*1e6} END {printf "sum %.2f", s}'; }
echo "== dot grids at 2^25 (min us, --only All)" | tee $O/dot_grids.txt
for cfg in "" "dot:grid=-3" "dot:grid=-6" "dot:grid=-9" "dot:grid=-12" "dot:grid=-15" "dot:block=256,grid=-6" "dot:block=256,grid=-12" "dot:block=256,grid=-18"; do
  echo "2^25 [$cfg]: $(BABELSTREAM_B200_TUNE="$cfg" drv build/b200-stream 33554432 All 100)" | tee -a $O/dot_grids.txt
done
echo "== fp32 rows of the per-size table" | tee $O/size_table_f32.txt
for s in 20 22 24 26 27; do
  n=$((1 << s))
  echo "f32 2^$s b200:   $(drv build/b200-stream $n All 40 --float)" | tee -a $O/size_table_f32.txt
  echo "f32 2^$s thrust: $(drv oracle/_ref/thrust-stream $n All 40 --float)" | tee -a $O/size_table_f32.txt
  echo "f32 2^$s cuda:   $(drv oracle/_ref/cuda-stream $n All 40 --float)" | tee -a $O/size_table_f32.txt
done
for s in 28 29; do
  n=$((1 << s))
  echo "f64 2^$s b200:   $(drv build/b200-stream $n All 40)" | tee -a $O/size_table_f32.txt
  echo "f64 2^$s thrust: $(drv oracle/_ref/thrust-stream $n All 40)" | tee -a $O/size_table_f32.txt
done
echo "== bench"
python bench.py --impl reference > $O/bench_reference.json 2> $O/bench_reference.err; echo "bench ref rc=$?"
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"; cut -c1-700 $O/bench_n1.json; tail -3 $O/bench_n1.err
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/smoke.log
echo "gpu_r2_i done"
