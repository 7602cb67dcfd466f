#!/bin/bash
# Round 2, GPU session M (1 GPU): --order Isolated (src/main.cpp:168-175: every kernel num_times in a row) for the three
# models, and --only Triad at 2^30.
set -u
O=gpurun_out/r2m
mkdir -p $O
drv() { "$1" --csv --gigabytes -s $2 --only $3 -n $4 "${@:5}" 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | awk -F, '{printf "%s %.2f  ", $1, $2*1e6; s+=$2*1e6} END {printf "sum %.2f", s}'; }
echo "== --order Isolated, min us" | tee $O/isolated.txt
for s in 22 25 29; do
  n=$((1 << s)); it=100; [ $s -ge 29 ] && it=20
  echo "2^$s b200 alt:    $(drv build/b200-stream $n All $it --order Isolated)" | tee -a $O/isolated.txt
  echo "2^$s b200 fwd:    $(BABELSTREAM_B200_SWEEP=forward drv build/b200-stream $n All $it --order Isolated)" | tee -a $O/isolated.txt
  echo "2^$s thrust:      $(drv oracle/_ref/thrust-stream $n All $it --order Isolated)" | tee -a $O/isolated.txt
  echo "2^$s cuda:        $(drv oracle/_ref/cuda-stream $n All $it --order Isolated)" | tee -a $O/isolated.txt
done
n=$((1 << 30))
echo "== --only Triad at 2^30" | tee -a $O/isolated.txt
for i in 1 2; do
  echo "2^30 b200:   $(drv build/b200-stream $n Triad 30)" | tee -a $O/isolated.txt
  echo "2^30 thrust: $(drv oracle/_ref/thrust-stream $n Triad 30)" | tee -a $O/isolated.txt
done
echo "2^30 b200 hint=3: $(BABELSTREAM_B200_TUNE="all:hint=3" drv build/b200-stream $n Triad 30)" | tee -a $O/isolated.txt
echo "2^30 b200 hint=1: $(BABELSTREAM_B200_TUNE="all:hint=1" drv build/b200-stream $n Triad 30)" | tee -a $O/isolated.txt
echo "2^30 b200 fwd hint=3: $(BABELSTREAM_B200_SWEEP=forward BABELSTREAM_B200_TUNE="all:hint=3" drv build/b200-stream $n Triad 30)" | tee -a $O/isolated.txt
echo "gpu_r2_m done"
