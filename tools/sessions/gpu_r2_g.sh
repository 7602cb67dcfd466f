#!/bin/bash
# Round 2, GPU session G (1 GPU): shortened elementwise prologue (affine sweep map) A/B against the previous commit's
# library, and cache-hint classes for the alternating sweep at the large sizes.
set -u
O=gpurun_out/r2g
mkdir -p $O
drv() { "$1" --csv --gigabytes -s $2 --only $3 -n $4 "${@:5}" 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | awk -F, '{printf "%s %.2f  ", $1, $2*1e6; s+=$2*1e6} END {printf "sum %.2f", s}'; }
NEW=build/b200-stream; OLD=build/old/build/b200-stream
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x > $O/pytest_kernels.log 2>&1; echo "pytest kernels rc=$?"; tail -3 $O/pytest_kernels.log
n=$((1 << 29))
for i in 1 2; do
  echo "2^29 old fwd Triad: $(drv $OLD $n Triad 60)" | tee -a $O/ab.txt
  echo "2^29 new fwd Triad: $(BABELSTREAM_B200_SWEEP=forward drv $NEW $n Triad 60)" | tee -a $O/ab.txt
  echo "2^29 new alt Triad: $(drv $NEW $n Triad 60)" | tee -a $O/ab.txt
  echo "2^29 new alt Triad hint=1: $(BABELSTREAM_B200_TUNE="all:hint=1" drv $NEW $n Triad 60)" | tee -a $O/ab.txt
  echo "2^29 new alt Triad hint=2: $(BABELSTREAM_B200_TUNE="all:hint=2" drv $NEW $n Triad 60)" | tee -a $O/ab.txt
  echo "2^29 new fwd Triad hint=1: $(BABELSTREAM_B200_SWEEP=forward BABELSTREAM_B200_TUNE="all:hint=1" drv $NEW $n Triad 60)" | tee -a $O/ab.txt
done
for cfg in "triad:hint=1,block=256" "triad:hint=1,block=1024" "triad:hint=1,unroll=2" "triad:hint=1,vec=32,block=256" "triad:hint=5" "triad:hint=4" "triad:hint=1,vec=32,block=512"; do
  echo "2^29 new alt [$cfg]: $(BABELSTREAM_B200_TUNE="$cfg" drv $NEW $n Triad 40)" | tee -a $O/ab.txt
done
for s in 29 28 27 26 25; do
  n=$((1 << s)); it=30; [ $s -le 26 ] && it=100
  for cfg in "" "all:hint=1" "all:hint=2" "all:hint=1;dot:hint=2" "all:hint=2;dot:hint=1"; do
    echo "2^$s new alt [$cfg]: $(BABELSTREAM_B200_TUNE="$cfg" drv $NEW $n All $it)" | tee -a $O/ab.txt
  done
  echo "2^$s new fwd []: $(BABELSTREAM_B200_SWEEP=forward drv $NEW $n All $it)" | tee -a $O/ab.txt
  echo "2^$s old: $(drv $OLD $n All $it)" | tee -a $O/ab.txt
  echo "2^$s thrust: $(drv oracle/_ref/thrust-stream $n All $it)" | tee -a $O/ab.txt
done
echo "gpu_r2_g done"
