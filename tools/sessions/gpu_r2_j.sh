#!/bin/bash
# Round 2, GPU session J (1 GPU): the 2^30 fp64 point (config 3's largest size), where thrust's mul/add/triad are ahead.
set -u
O=gpurun_out/r2j
mkdir -p $O
drv() { "$1" --csv --gigabytes -s $2 --only $3 -n $4 "${@:5}" 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | awk -F, '{printf "%s %.2f  ", $1, $2*1e6; s+=$2*1e6} END {printf "sum %.2f", s}'; }
timeout 500 python tools/sweep.py --kernels mul,add,triad --log2n 30 --blocking --alternate --quick --hints 1,2,3 --top 12 --out $O/sweep_alt_2p30.csv > $O/sweep_alt_2p30.log 2>&1; grep -A7 "^==" $O/sweep_alt_2p30.log
n=$((1 << 30))
echo "== driver at 2^30" | tee $O/driver_2p30.txt
for cfg in "alternate|" "forward|" "alternate|all:hint=1" "alternate|all:hint=3" "forward|all:hint=3" "alternate|all:hint=4"; do
  sw=${cfg%%|*}; tn=${cfg#*|}
  echo "2^30 b200 [$sw $tn]: $(BABELSTREAM_B200_SWEEP=$sw BABELSTREAM_B200_TUNE="$tn" drv build/b200-stream $n All 12)" | tee -a $O/driver_2p30.txt
done
echo "2^30 thrust: $(drv oracle/_ref/thrust-stream $n All 12)" | tee -a $O/driver_2p30.txt
echo "2^30 thrust: $(drv oracle/_ref/thrust-stream $n All 12)" | tee -a $O/driver_2p30.txt
echo "2^30 b200 []: $(drv build/b200-stream $n All 12)" | tee -a $O/driver_2p30.txt
echo "gpu_r2_j done"
