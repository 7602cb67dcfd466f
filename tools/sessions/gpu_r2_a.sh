#!/bin/bash
# Round 2, GPU session A (1 GPU): what runs in the reference's thrust model (ncu), the ew_lab sweeps of
# structurally different elementwise candidates, launch/wait latency flavours, same-box driver baselines.
set -u
O=gpurun_out/r2a
mkdir -p $O
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > $O/clocks.csv &
SMI=$!
drv() { # binary size only n
  "$1" --csv --gigabytes -s $2 --only $3 -n $4 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,5,6 | tr '\n' ' '
}
echo "== driver baselines (GB/s,min s)" | tee $O/driver_baselines.txt
for rep in 1 2; do
  for b in build/b200-stream oracle/_ref/cuda-stream oracle/_ref/thrust-stream; do
    echo "2^25 All $b: $(drv $b 33554432 All 100)" | tee -a $O/driver_baselines.txt
  done
done
for b in build/b200-stream oracle/_ref/cuda-stream oracle/_ref/thrust-stream; do
  echo "2^29 Triad $b: $(drv $b 536870912 Triad 30)" | tee -a $O/driver_baselines.txt
done
for s in 1048576 2097152 4194304 8388608 16777216; do
  for b in build/b200-stream oracle/_ref/cuda-stream oracle/_ref/thrust-stream; do
    echo "s=$s All $b: $(drv $b $s All 200)" | tee -a $O/driver_baselines.txt
  done
done

echo "== lab"
./build/ew_lab lat > $O/lat.txt 2>&1; cat $O/lat.txt
for l in 20 21 22; do ./build/ew_lab alt $l; done > $O/alt.txt 2>&1; cat $O/alt.txt
for op in triad add copy mul nstream; do
  timeout 240 ./build/ew_lab b2b 29 $op 1 > $O/b2b_29_$op.csv 2> $O/b2b_29_$op.err; echo "b2b 29 $op rc=$?"
  timeout 120 ./build/ew_lab b2b 25 $op 1 > $O/b2b_25_$op.csv 2> $O/b2b_25_$op.err; echo "b2b 25 $op rc=$?"
  timeout 120 ./build/ew_lab seq 25 $op 1 > $O/seq_25_$op.csv 2> $O/seq_25_$op.err; echo "seq 25 $op rc=$?"
done
for f in $O/b2b_29_triad.csv $O/seq_25_triad.csv $O/seq_25_add.csv; do echo "-- top of $f"; sort -t, -k4 -g -r $f | head -12; done
grep -l -v "^$" $O/*.err 2>/dev/null | head; cat $O/*.err | head -20

echo "== ncu: the reference's thrust model"
ncu_one() { # name regex skip count cmd...
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 4
  "$@" > $O/plain_$name.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k "regex:$rx" -s $skip -c $cnt -f -o $O/prof_$name "$@" > $O/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
  ncu -i $O/prof_$name.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__cycles_active.avg.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,launch__occupancy_limit_registers,launch__waves_per_multiprocessor,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum > $O/summary_$name.csv 2>&1
}
ncu_one thrust_all_2p25 "transform_kernel|DeviceReduce" 0 14 ./oracle/_ref/thrust-stream -s 33554432 --only All -n 3
ncu_one thrust_triad_2p29 "transform_kernel" 1 2 ./oracle/_ref/thrust-stream -s 536870912 --only Triad -n 4
ncu_one b200_all_2p25 "ew_kernel|dot_kernel" 8 6 ./build/b200-stream -s 33554432 --only All -n 3
kill $SMI
echo "gpu_r2_a done"
