#!/bin/bash
# Round 2, GPU session N (1 GPU): first-call costs with and without b200_warmup, --order Isolated again.
set -u
O=gpurun_out/r2n
mkdir -p $O
drv() { "$1" --csv --gigabytes -s $2 --only $3 -n $4 "${@:5}" 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | awk -F, '{printf "%s %.2f  ", $1, $2*1e6; s+=$2*1e6} END {printf "sum %.2f", s}'; }
python tools/first_call_probe.py 22 cold 2>&1 | tee $O/first_call.txt
python tools/first_call_probe.py 22 warm 2>&1 | tee -a $O/first_call.txt
python tools/first_call_probe.py 25 warm 2>&1 | tee -a $O/first_call.txt
echo "== --order Isolated, min us (after the constructor warm-up)" | tee $O/isolated.txt
for s in 20 22 25 29; do
  n=$((1 << s)); it=100; [ $s -ge 29 ] && it=20
  echo "2^$s b200 alt:    $(drv build/b200-stream $n All $it --order Isolated)" | tee -a $O/isolated.txt
  echo "2^$s b200 fwd:    $(BABELSTREAM_B200_SWEEP=forward drv build/b200-stream $n All $it --order Isolated)" | tee -a $O/isolated.txt
  echo "2^$s thrust:      $(drv oracle/_ref/thrust-stream $n All $it --order Isolated)" | tee -a $O/isolated.txt
  echo "2^$s cuda:        $(drv oracle/_ref/cuda-stream $n All $it --order Isolated)" | tee -a $O/isolated.txt
done
echo "2^25 b200 classic: $(drv build/b200-stream 33554432 All 100)" | tee -a $O/isolated.txt
timeout 900 python -m pytest tests/test_stream_gpu.py -m gpu -q -x 2>&1 | tail -3
echo "gpu_r2_n done"
