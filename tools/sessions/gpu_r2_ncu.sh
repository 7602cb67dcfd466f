#!/bin/bash
# Round 2, ncu session (1 GPU, ncu is the only tool of this call): launch list of the bench command, one --set full
# capture of the six kernels inside the timed region (2^25) and of triad at 2^29, summaries as CSV.
set -u
O=gpurun_out/r2ncu
mkdir -p $O
BENCH="python bench.py --steps 2 --warmup 3 --no-extras"
$BENCH > $O/bench_small.json 2> $O/bench_small.err; echo "bench (no ncu) rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench_cmd.csv $BENCH > $O/ncu_list.log 2>&1
echo "ncu list rc=$?"; grep -c "ew_kernel\|dot_kernel" $O/launches_bench_cmd.csv
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__cycles_active.avg.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,launch__waves_per_multiprocessor,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum,launch__occupancy_limit_warps
ncu_one() { # name regex skip count cmd...
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 4
  "$@" > /dev/null 2>&1; echo "$name without ncu rc=$?"
  ncu --set full --clock-control none -k "regex:$rx" -s $skip -c $cnt -f -o $O/prof_$name "$@" > $O/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
  ncu -i $O/prof_$name.ncu-rep --page raw --csv --metrics $M > $O/summary_$name.csv 2>&1
  ls -la $O/prof_$name.ncu-rep | awk '{print $5}'
}
# the driver's own sequence: iteration 3 of --only All = copy mul add triad nstream dot (12 kernels skipped)
ncu_one b200_all_2p25 "ew_kernel|dot_kernel" 13 6 ./build/b200-stream -s 33554432 --only All -n 4
ncu_one b200_triad_2p29 "ew_kernel" 2 2 ./build/b200-stream -s 536870912 --only Triad -n 5
ncu_one b200_all_2p20 "ew_kernel|dot_kernel" 13 6 ./build/b200-stream -s 1048576 --only All -n 4
ncu_one cuda_all_2p20 "_kernel" 13 6 ./oracle/_ref/cuda-stream -s 1048576 --only All -n 4
# dot alone (no predecessor's dirty lines): --only Dot
ncu_one b200_dot_alone_2p25 "dot_kernel" 2 2 ./build/b200-stream -s 33554432 --only Dot -n 5
cat $O/summary_b200_all_2p25.csv | cut -c1-400
du -sh gpurun_out; du -s gpurun_out | awk '{ if ($1 > 50000) print "PRUNE" }' | grep -q PRUNE && rm -f $O/*.ncu-rep
echo "gpu_r2_ncu done"
