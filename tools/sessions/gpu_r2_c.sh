#!/bin/bash
# Round 2, GPU session C (1 GPU): parity suite on the rebuilt library, driver comparisons, class calibration, bench.
set -u
O=gpurun_out/r2c
mkdir -p $O
drv() { # binary size only n  [env...]
  "$1" --csv --gigabytes -s $2 --only $3 -n $4 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | awk -F, '{printf "%s %.2f  ", $1, $2*1e6}'
}
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 $O/pytest_gpu.log

echo "== 2^25 All, min call time in us" | tee $O/driver.txt
for rep in 1 2; do
  echo "b200 graph : $(drv build/b200-stream 33554432 All 100)" | tee -a $O/driver.txt
  echo "b200 direct: $(BABELSTREAM_B200_LAUNCH=direct drv build/b200-stream 33554432 All 100)" | tee -a $O/driver.txt
  echo "b200 cmake : $(drv build/cmake/b200-stream 33554432 All 100)" | tee -a $O/driver.txt
  echo "cuda-stream: $(drv oracle/_ref/cuda-stream 33554432 All 100)" | tee -a $O/driver.txt
  echo "thrust     : $(drv oracle/_ref/thrust-stream 33554432 All 100)" | tee -a $O/driver.txt
done
echo "== 2^29 Triad" | tee -a $O/driver.txt
for rep in 1 2 3; do echo "b200 graph : $(drv build/b200-stream 536870912 Triad 30)" | tee -a $O/driver.txt; done
echo "b200 direct: $(BABELSTREAM_B200_LAUNCH=direct drv build/b200-stream 536870912 Triad 30)" | tee -a $O/driver.txt
echo "b200 16x2x256: $(BABELSTREAM_B200_TUNE='triad:vec=16,unroll=2,block=256' drv build/b200-stream 536870912 Triad 30)" | tee -a $O/driver.txt
echo "thrust     : $(drv oracle/_ref/thrust-stream 536870912 Triad 30)" | tee -a $O/driver.txt
echo "cuda-stream: $(drv oracle/_ref/cuda-stream 536870912 Triad 30)" | tee -a $O/driver.txt

echo "== size classes (min us per call)" | tee $O/classes.txt
for s in 1048576 2097152 4194304; do
  echo "s=$s cuda-stream : $(drv oracle/_ref/cuda-stream $s All 300)" | tee -a $O/classes.txt
  for cfg in "" "all:vec=16,unroll=4,block=128,hint=1" "all:vec=32,unroll=1,block=256,hint=2" "all:vec=16,unroll=2,block=256,hint=1" "all:vec=32,unroll=2,block=128,hint=1"; do
    echo "s=$s b200 [$cfg]: $(BABELSTREAM_B200_TUNE="$cfg" drv build/b200-stream $s All 300)" | tee -a $O/classes.txt
  done
done
for s in 8388608 16777216; do
  echo "s=$s cuda-stream : $(drv oracle/_ref/cuda-stream $s All 200)" | tee -a $O/classes.txt
  echo "s=$s thrust      : $(drv oracle/_ref/thrust-stream $s All 200)" | tee -a $O/classes.txt
  for cfg in "" "all:hint=1" "all:hint=3" "all:vec=16,unroll=1,block=256,hint=2" "all:vec=16,unroll=1,block=256,hint=3" "all:vec=16,unroll=2,block=512,hint=2" "all:vec=32,unroll=1,block=256,hint=2" "all:vec=32,unroll=1,block=256,hint=1" "all:vec=16,unroll=1,block=128,hint=2"; do
    echo "s=$s b200 [$cfg]: $(BABELSTREAM_B200_TUNE="$cfg" drv build/b200-stream $s All 200)" | tee -a $O/classes.txt
  done
done
for s in 33554432 67108864 134217728 268435456; do
  for cfg in "" "all:vec=16,unroll=1,block=256" "all:vec=16,unroll=1,block=512" "all:vec=16,unroll=2,block=256" "all:vec=16,unroll=2,block=512" "all:vec=32,unroll=1,block=256" "all:vec=16,unroll=1,block=128"; do
    echo "s=$s b200 [$cfg]: $(BABELSTREAM_B200_TUNE="$cfg" drv build/b200-stream $s All 60)" | tee -a $O/classes.txt
  done
done
echo "== fp32"
for s in 67108864; do
  echo "s=$s float cuda-stream: $(./oracle/_ref/cuda-stream --csv --gigabytes -s $s --float -n 100 2>/dev/null | grep -E '^(Copy|Mul|Add|Triad|Dot),' | cut -d, -f1,5 | tr '\n' ' ')" | tee -a $O/driver.txt
  echo "s=$s float b200       : $(./build/b200-stream --csv --gigabytes -s $s --float -n 100 2>/dev/null | grep -E '^(Copy|Mul|Add|Triad|Dot),' | cut -d, -f1,5 | tr '\n' ' ')" | tee -a $O/driver.txt
done
echo "== bench"
timeout 600 python bench.py --impl reference --steps 30 --warmup 3 > $O/bench_reference.json 2> $O/bench_reference.err; echo "bench ref rc=$?"; cut -c1-600 $O/bench_reference.json
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; cat $O/bench.json; tail -5 $O/bench.err
echo "gpu_r2_c done"
