#!/bin/bash
# Round 2, GPU session B (1 GPU): ew_lab sweeps (all ops), launch/completion paths on the real kernel,
# PDL / graph for the six-kernel step, ncu on the reference's thrust model.  Keeps gpurun_out small.
set -u
O=gpurun_out/r2b
mkdir -p $O
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 1000 > $O/clocks.csv &
SMI=$!
echo "== path"
for l in 20 25 29; do ./build/ew_lab path $l triad; done > $O/path.txt 2>&1; cat $O/path.txt
./build/ew_lab path 25 add >> $O/path.txt 2>&1
echo "== seq6"
./build/ew_lab seq6 25 > $O/seq6.txt 2>&1; cat $O/seq6.txt
echo "== sweeps"
for op in triad add copy mul nstream; do
  timeout 200 ./build/ew_lab seq 25 $op 1 > $O/seq_25_$op.csv 2> $O/seq_25_$op.err; echo "seq 25 $op rc=$?"
  timeout 200 ./build/ew_lab b2b 25 $op 1 > $O/b2b_25_$op.csv 2> $O/b2b_25_$op.err; echo "b2b 25 $op rc=$?"
  timeout 300 ./build/ew_lab b2b 29 $op 1 > $O/b2b_29_$op.csv 2> $O/b2b_29_$op.err; echo "b2b 29 $op rc=$?"
done
timeout 300 ./build/ew_lab seq 29 triad 1 > $O/seq_29_triad.csv 2> $O/seq_29_triad.err; echo "seq 29 triad rc=$?"
for f in $O/seq_25_*.csv $O/seq_29_triad.csv; do echo "-- $f: baseline, thrust-like, top 6"; sed -n 2,3p $f; grep -v REFUSED $f | sort -t, -k4 -g -r | head -6; done
grep -h REFUSED $O/*.csv | head -5
cat $O/*.err | head -5

echo "== ncu: the reference's thrust model (summaries only; reports pruned to stay under the copy-back limit)"
ncu_one() { # name regex skip count cmd...
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 4
  "$@" > $O/plain_$name.log 2>&1 &&
  ncu --set full --clock-control none -k "regex:$rx" -s $skip -c $cnt -f -o $O/prof_$name "$@" > $O/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
  ncu -i $O/prof_$name.ncu-rep --page raw --csv > $O/raw_$name.csv 2>&1
  ls -la $O/prof_$name.ncu-rep
}
ncu_one thrust_ew_2p25 "transform_kernel" 5 5 ./oracle/_ref/thrust-stream -s 33554432 --only All -n 3
ncu_one thrust_triad_2p29 "transform_kernel" 1 1 ./oracle/_ref/thrust-stream -s 536870912 --only Triad -n 3
kill $SMI
du -sh gpurun_out; du -s gpurun_out | awk '{ if ($1 > 55000) print "PRUNE" }' | grep -q PRUNE && rm -f $O/*.ncu-rep
du -sh gpurun_out
echo "gpu_r2_b done"
