#!/bin/bash
# Round 3, ncu session (1 GPU): launch list of the bench command; --set full summaries of the six kernels at 2^25 AND 2^29,
# of the bulk / TMA copy variants, of dot alone; and one LIVE six-kernel sequence without cache control (application replay)
# that shows the alternating sweep's L2 hand-over.
set -u
O=gpurun_out/r3ncu
mkdir -p $O
BENCH="python bench.py --steps 2 --warmup 3 --no-extras"
$BENCH > $O/bench_small.json 2> $O/bench_small.err; echo "bench (no ncu) rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench_cmd.csv $BENCH > $O/ncu_list.log 2>&1
echo "ncu list rc=$?"; grep -c "ew_kernel\|dot_kernel" $O/launches_bench_cmd.csv
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__cycles_active.avg.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,launch__waves_per_multiprocessor,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum,launch__occupancy_limit_warps
ncu_one() { # name regex skip count cmd...
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 4
  "$@" > /dev/null 2>&1; echo "$name without ncu rc=$?"
  timeout 600 ncu --set full --clock-control none -k "regex:$rx" -s $skip -c $cnt -f -o $O/prof_$name "$@" > $O/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
  ncu -i $O/prof_$name.ncu-rep --page raw --csv --metrics $M > $O/summary_$name.csv 2>&1
  rm -f $O/prof_$name.ncu-rep
}
# launches of one driver run: 12 warm-up kernels (constructor), init, then copy mul add triad dot nstream per iteration
ncu_one b200_all_2p25 "ew_kernel|dot_kernel" 31 6 ./build/b200-stream -s 33554432 --only All -n 6
ncu_one b200_all_2p29 "ew_kernel|dot_kernel" 31 6 ./build/b200-stream -s 536870912 --only All -n 6
ncu_one b200_dot_alone_2p25 "dot_kernel" 4 2 ./build/b200-stream -s 33554432 --only Dot -n 6
export BABELSTREAM_B200_COPY_VARIANT=bulk
ncu_one b200_copy_bulk_2p25 "copy_bulk_kernel" 3 2 ./build/b200-stream -s 33554432 --only Copy -n 6
ncu_one b200_copy_bulk_2p29 "copy_bulk_kernel" 3 2 ./build/b200-stream -s 536870912 --only Copy -n 6
export BABELSTREAM_B200_COPY_VARIANT=tma
ncu_one b200_copy_tma_2p25 "copy_tma_kernel" 3 2 ./build/b200-stream -s 33554432 --only Copy -n 6
ncu_one b200_copy_tma_2p29 "copy_tma_kernel" 3 2 ./build/b200-stream -s 536870912 --only Copy -n 6
unset BABELSTREAM_B200_COPY_VARIANT
# the live sequence: no cache flush between kernels, each pass is a fresh run of the program (application replay)
L=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,dram__cycles_active.avg.pct_of_peak_sustained_elapsed
for sw in alternate forward; do
  BABELSTREAM_B200_SWEEP=$sw timeout 600 ncu --replay-mode application --cache-control none --clock-control none --metrics $L \
     -k "regex:ew_kernel|dot_kernel" -s 31 -c 12 --csv --log-file $O/live_sequence_2p25_$sw.csv ./build/b200-stream -s 33554432 --only All -n 6 > $O/ncu_live_$sw.log 2>&1
  echo "live $sw rc=$?"
done
cut -c1-300 $O/summary_b200_all_2p29.csv | tail -8
tail -n 14 $O/live_sequence_2p25_alternate.csv | cut -c1-260
du -sh gpurun_out
echo "gpu_r3_ncu done"
