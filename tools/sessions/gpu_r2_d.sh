#!/bin/bash
# Round 2, GPU session D (1 GPU): parity suite, zigzag (alternating sweep direction + L2 keep) experiment, dot grids.
set -u
O=gpurun_out/r2d
mkdir -p $O
drv() { "$1" --csv --gigabytes -s $2 --only $3 -n $4 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | awk -F, '{printf "%s %.2f  ", $1, $2*1e6}'; }
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 $O/pytest_gpu.log
echo "== zigzag"
./build/ew_lab zz 25 > $O/zz.txt 2>&1; ./build/ew_lab zz 29 1 >> $O/zz.txt 2>&1; ./build/ew_lab zz 27 >> $O/zz.txt 2>&1; ./build/ew_lab zz 23 >> $O/zz.txt 2>&1; cat $O/zz.txt
echo "== dot grids at 2^25 and 2^29 (--only All, us)" | tee $O/dot.txt
for cfg in "" "dot:grid=-12" "dot:grid=-16" "dot:grid=-27" "dot:grid=-55" "dot:grid=-110" "dot:block=256,grid=-16" "dot:block=256,grid=-32" "dot:block=256,grid=-55" "dot:block=256,grid=-110" "dot:block=256,unroll=4,grid=-55" "dot:block=256,unroll=1,grid=-110" "dot:block=128,unroll=4,grid=-110" "dot:vec=32,unroll=1,block=256,grid=-55" "dot:hint=3" "dot:hint=3,grid=-27" "dot:hint=1,grid=-27"; do
  echo "2^25 [$cfg]: $(BABELSTREAM_B200_TUNE="$cfg" drv build/b200-stream 33554432 All 100)" | tee -a $O/dot.txt
done
for cfg in "" "dot:grid=-27" "dot:block=256,grid=-55" "dot:block=256,grid=-110"; do
  echo "2^29 [$cfg]: $(BABELSTREAM_B200_TUNE="$cfg" drv build/b200-stream 536870912 All 20)" | tee -a $O/dot.txt
done
for cfg in "" "dot:grid=-27" "dot:block=256,grid=-55"; do
  echo "2^23 [$cfg]: $(BABELSTREAM_B200_TUNE="$cfg" drv build/b200-stream 8388608 All 200)" | tee -a $O/dot.txt
done
echo "gpu_r2_d done"
