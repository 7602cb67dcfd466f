#!/bin/bash
# Round 2, GPU session F (1 GPU): same-box A/B of the previous commit's library (build/old, a git worktree of c0e2b5a)
# against the current one at the large sizes, plus launch-shape variants of the large-size classes.
set -u
O=gpurun_out/r2f
mkdir -p $O
drv() { "$1" --csv --gigabytes -s $2 --only $3 -n $4 "${@:5}" 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | awk -F, '{printf "%s %.2f  ", $1, $2*1e6; s+=$2*1e6} END {printf "sum %.2f", s}'; }
NEW=build/b200-stream; OLD=build/old/build/b200-stream
n=$((1 << 29))
for i in 1 2 3; do
  echo "2^29 old fwd Triad: $(drv $OLD $n Triad 60)" | tee -a $O/ab.txt
  echo "2^29 new fwd Triad: $(BABELSTREAM_B200_SWEEP=forward drv $NEW $n Triad 60)" | tee -a $O/ab.txt
  echo "2^29 new alt Triad: $(drv $NEW $n Triad 60)" | tee -a $O/ab.txt
  echo "2^29 thrust  Triad: $(drv oracle/_ref/thrust-stream $n Triad 60)" | tee -a $O/ab.txt
done
for cfg in "triad:hint=2" "triad:hint=1" "triad:block=256" "triad:block=1024" "triad:vec=32,block=256" "triad:unroll=2,block=256" "triad:unroll=2" "triad:grid=-4" "triad:grid=-8" "triad:vec=8,unroll=2,block=128" "triad:vec=8,unroll=2,block=128,hint=1" "triad:vec=8,unroll=2,block=128,hint=5" "triad:vec=16,block=128" "triad:vec=16,block=128,hint=1"; do
  echo "2^29 new alt [$cfg]: $(BABELSTREAM_B200_TUNE="$cfg" drv $NEW $n Triad 40)" | tee -a $O/ab.txt
done
n=$((1 << 28))
for i in 1 2; do
  echo "2^28 old All: $(drv $OLD $n All 40)" | tee -a $O/ab.txt
  echo "2^28 new fwd All: $(BABELSTREAM_B200_SWEEP=forward drv $NEW $n All 40)" | tee -a $O/ab.txt
  echo "2^28 new alt All: $(drv $NEW $n All 40)" | tee -a $O/ab.txt
  echo "2^28 thrust All: $(drv oracle/_ref/thrust-stream $n All 40)" | tee -a $O/ab.txt
done
for cfg in "add:hint=2;mul:hint=2" "add:hint=1;mul:hint=1" "add:block=256;mul:block=256" "add:vec=8,unroll=2,block=128;mul:vec=8,unroll=2,block=128" "add:vec=8,unroll=2,block=128,hint=1;mul:vec=8,unroll=2,block=128,hint=1" "add:vec=16,block=128;mul:vec=16,block=128" "add:vec=32,block=256;mul:vec=32,block=256" "add:unroll=2;mul:unroll=1" "add:grid=-4;mul:grid=-4" "add:grid=-16;mul:grid=-16"; do
  echo "2^28 new alt [$cfg]: $(BABELSTREAM_B200_TUNE="$cfg" drv $NEW $n All 40)" | tee -a $O/ab.txt
done
n=$((1 << 25))
for i in 1 2; do
  echo "2^25 old All: $(drv $OLD $n All 100)" | tee -a $O/ab.txt
  echo "2^25 new fwd All: $(BABELSTREAM_B200_SWEEP=forward drv $NEW $n All 100)" | tee -a $O/ab.txt
  echo "2^25 new alt keep=128: $(BABELSTREAM_B200_TUNE="all:keep=128" drv $NEW $n All 100)" | tee -a $O/ab.txt
  echo "2^25 new alt keep=160: $(BABELSTREAM_B200_TUNE="all:keep=160" drv $NEW $n All 100)" | tee -a $O/ab.txt
done
echo "gpu_r2_f done"
