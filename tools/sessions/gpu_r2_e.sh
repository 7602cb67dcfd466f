#!/bin/bash
# Round 2, GPU session E (1 GPU): parity suite with the alternating sweep in the product, driver-timed comparison of
# sweep/keep/dot-grid variants against the reference's cuda and thrust models, bench.py (both arms).
set -u
O=gpurun_out/r2e
mkdir -p $O
drv() { "$1" --csv --gigabytes -s $2 --only $3 -n $4 "${@:5}" 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | awk -F, '{printf "%s %.2f  ", $1, $2*1e6; s+=$2*1e6} END {printf "sum %.2f", s}'; }
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_gpu.log
echo "== driver, --only All (Copy Mul Add Triad Dot) and Nstream, min us" | tee $O/driver.txt
for s in 25; do
  n=$((1 << s))
  for cfg in "sweep=alternate|" "sweep=forward|" "sweep=alternate|all:keep=-1" "sweep=alternate|all:keep=48" "sweep=alternate|all:keep=64" "sweep=alternate|all:keep=128" "sweep=alternate|all:keep=192" \
             "sweep=alternate|dot:grid=-16" "sweep=alternate|dot:grid=-27" "sweep=alternate|dot:grid=-4" "sweep=alternate|dot:flags=4" "sweep=alternate|dot:hint=3" "sweep=alternate|dot:hint=1" "sweep=alternate|dot:keep=-1"; do
    sw=${cfg%%|*}; tn=${cfg#*|}
    echo "2^$s b200 [$sw $tn]: $(BABELSTREAM_B200_SWEEP=${sw#sweep=} BABELSTREAM_B200_TUNE="$tn" drv build/b200-stream $n All 100)" | tee -a $O/driver.txt
  done
  echo "2^$s b200 nstream alt: $(drv build/b200-stream $n Nstream 100)   fwd: $(BABELSTREAM_B200_SWEEP=forward drv build/b200-stream $n Nstream 100)" | tee -a $O/driver.txt
  echo "2^$s thrust:  $(drv oracle/_ref/thrust-stream $n All 100)" | tee -a $O/driver.txt
  echo "2^$s cuda:    $(drv oracle/_ref/cuda-stream $n All 100)" | tee -a $O/driver.txt
  echo "2^$s b200 again: $(drv build/b200-stream $n All 100)" | tee -a $O/driver.txt
done
for s in 20 22 23 24 26 27 28; do
  n=$((1 << s)); it=100; [ $s -ge 27 ] && it=40
  echo "2^$s b200 alt: $(drv build/b200-stream $n All $it)" | tee -a $O/driver.txt
  echo "2^$s b200 fwd: $(BABELSTREAM_B200_SWEEP=forward drv build/b200-stream $n All $it)" | tee -a $O/driver.txt
  echo "2^$s b200 alt keep=-1: $(BABELSTREAM_B200_TUNE="all:keep=-1" drv build/b200-stream $n All $it)" | tee -a $O/driver.txt
  echo "2^$s thrust:   $(drv oracle/_ref/thrust-stream $n All $it)" | tee -a $O/driver.txt
  echo "2^$s cuda:     $(drv oracle/_ref/cuda-stream $n All $it)" | tee -a $O/driver.txt
done
n=$((1 << 29))
for i in 1 2; do
  echo "2^29 b200 alt Triad: $(drv build/b200-stream $n Triad 100)" | tee -a $O/driver.txt
  echo "2^29 b200 fwd Triad: $(BABELSTREAM_B200_SWEEP=forward drv build/b200-stream $n Triad 100)" | tee -a $O/driver.txt
done
echo "2^29 b200 alt All: $(drv build/b200-stream $n All 30)" | tee -a $O/driver.txt
echo "2^29 b200 fwd All: $(BABELSTREAM_B200_SWEEP=forward drv build/b200-stream $n All 30)" | tee -a $O/driver.txt
echo "2^29 thrust Triad: $(drv oracle/_ref/thrust-stream $n Triad 100)" | tee -a $O/driver.txt
echo "2^29 thrust All: $(drv oracle/_ref/thrust-stream $n All 30)" | tee -a $O/driver.txt
echo "2^29 cuda All: $(drv oracle/_ref/cuda-stream $n All 30)" | tee -a $O/driver.txt
echo "== fp32 2^26"
echo "b200: $(drv build/b200-stream $((1<<26)) All 100 --float)"
echo "== bench"
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"; cat $O/bench_n1.json; tail -3 $O/bench_n1.err
echo "gpu_r2_e done"
