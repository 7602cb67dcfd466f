#!/bin/bash
# One GPU-box session: parity suite, smoke, both bench arms, ncu evidence, tuning sweeps.
# Usage (under gpurun): bash tools/gpu_round.sh [tests|bench|ncu|sweep|driver ...]   default: all
set -u
mkdir -p gpurun_out
STAGES="${*:-tests bench ncu sweep driver}"
has() { [[ " $STAGES " == *" $1 "* ]]; }
BENCH_SMALL="python bench.py --steps 2 --warmup 3 --no-extras"

if has tests; then
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
  timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
fi
if has bench; then
  timeout 300 python bench.py --impl reference --steps 30 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench ref rc=$?"
  timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
fi
if has ncu; then
  $BENCH_SMALL > gpurun_out/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $BENCH_SMALL > gpurun_out/ncu_list.log 2>&1
  echo "ncu list rc=$?"
  $BENCH_SMALL > gpurun_out/plain2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k "regex:ew_kernel|dot_kernel" -s 20 -c 12 -f -o gpurun_out/prof_timed $BENCH_SMALL > gpurun_out/ncu_full.log 2>&1
  echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log
fi
if has sweep; then
  timeout 900 python tools/sweep.py --kernels triad,copy,mul,add,nstream,dot,init --log2n 29,25 --dtype f64 --out gpurun_out/sweep_f64.csv > gpurun_out/sweep_f64.log 2>&1; echo "sweep f64 rc=$?"
  timeout 600 python tools/sweep.py --quick --kernels triad,copy,mul,add,nstream,dot --log2n 30,26 --dtype f32 --out gpurun_out/sweep_f32.csv > gpurun_out/sweep_f32.log 2>&1; echo "sweep f32 rc=$?"
  timeout 300 python tools/sweep.py --kernels copy_async --log2n 29,25 --out gpurun_out/sweep_copy_async.csv > gpurun_out/sweep_copy_async.log 2>&1
  timeout 200 python tools/sweep.py --pad-test 29 --out gpurun_out/pad_test.csv > gpurun_out/pad_test.log 2>&1; echo "pad rc=$?"
fi
if has sweep2; then
  timeout 600 python tools/sweep.py --quick --kernels triad,add,copy,mul,nstream,init --log2n 29,25 --dtype f64 --out gpurun_out/sweep2_f64.csv > gpurun_out/sweep2_f64.log 2>&1; echo "sweep2 rc=$?"
  timeout 300 python tools/sweep.py --quick --kernels triad,copy,dot,nstream --log2n 20,22,23 --dtype f64 --out gpurun_out/sweep2_small.csv > gpurun_out/sweep2_small.log 2>&1
  grep -A4 "==" gpurun_out/sweep2_f64.log | head -80; grep -A3 "==" gpurun_out/sweep2_small.log | head -60
fi
if has sizes; then
  timeout 1500 python tools/size_sweep.py --out gpurun_out/size_sweep.csv > gpurun_out/size_sweep.log 2>&1; echo "sizes rc=$?"; cat gpurun_out/size_sweep.log
fi
if has block; then
  timeout 600 python tools/sweep.py --blocking --quick --kernels dot,triad,add,copy,mul,nstream --log2n 25,23 --dtype f64 --out gpurun_out/sweep_blocking.csv > gpurun_out/sweep_blocking.log 2>&1; echo "blocking sweep rc=$?"
  grep -A5 "==" gpurun_out/sweep_blocking.log | head -120
  timeout 300 python tools/sweep.py --blocking --kernels dot --log2n 25,27 --dtype f64 --out gpurun_out/sweep_blocking_dot.csv > gpurun_out/sweep_blocking_dot.log 2>&1
  grep -A8 "==" gpurun_out/sweep_blocking_dot.log | head -40
fi
if has smallcfg; then
  for s in 1048576 2097152 4194304; do
    ./oracle/_ref/cuda-stream --csv --gigabytes -s $s --only All -n 20 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,5 | tr '\n' ' ' | sed "s/^/s=$s ref_cuda: /"; echo
    for cfg in "" "all:vec=8,unroll=4,block=256" "all:vec=16,unroll=2,block=256" "all:vec=16,unroll=1,block=256" "all:vec=32,unroll=1,block=256" "all:vec=16,unroll=4,block=256" "all:vec=32,unroll=2,block=256" "all:vec=16,unroll=4,block=128" "all:vec=16,unroll=2,block=128;dot:block=256,grid=-4,unroll=4" "dot:block=256,grid=-2,unroll=4,vec=32" "dot:block=512,grid=-1,unroll=2,vec=32" "dot:block=256,grid=-8,unroll=4,vec=32"; do
      BABELSTREAM_B200_TUNE="$cfg" ./build/b200-stream --csv --gigabytes -s $s --only All -n 20 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,5 | tr '\n' ' ' | sed "s/^/s=$s [$cfg]: /"; echo
    done
  done
fi
if has small; then
  timeout 500 python tools/sweep.py --blocking --quick --kernels copy,mul,add,triad,nstream,dot,init --log2n 20,21,22 --dtype f64 --out gpurun_out/sweep_blocking_small.csv > gpurun_out/sweep_blocking_small.log 2>&1; echo "small rc=$?"
  grep -A4 "==" gpurun_out/sweep_blocking_small.log | head -150
fi
if has dotseq; then
  timeout 400 python tools/sweep.py --blocking --after triad --kernels dot --log2n 25,27,29 --dtype f64 --out gpurun_out/sweep_dot_after_triad.csv > gpurun_out/sweep_dot_after_triad.log 2>&1
  grep -A10 "==" gpurun_out/sweep_dot_after_triad.log | head -40
fi
if has overhead; then
  for i in 1 2; do
    ./oracle/_ref/cuda-stream --csv -s 1024 -n 3000 2>&1 | grep -E "^(Copy|Triad|Dot)," | cut -d, -f1,6,8
    BABELSTREAM_B200_HOSTPROF=1 ./build/b200-stream --csv -s 1024 -n 3000 2>&1 | grep -E "^(Copy|Triad|Dot),|hostprof" | cut -d, -f1,6,8
    BABELSTREAM_B200_SYNC=spin BABELSTREAM_B200_HOSTPROF=1 ./build/b200-stream --csv -s 1024 -n 3000 2>&1 | grep -E "^(Copy|Triad|Dot),|hostprof" | cut -d, -f1,6,8
  done
fi
if has waitmodes; then
  # blocking-call latency: cudaStreamSynchronize vs the kernel-rung doorbell, against the reference, driver-timed
  for s in 1024 1048576 4194304 33554432; do
    for rep in 1 2; do
      ./oracle/_ref/cuda-stream --csv --gigabytes -s $s --only All -n 300 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | tr '\n' ' ' | sed "s/^/s=$s ref_cuda        : /"; echo
      for m in stream doorbell auto; do
        BABELSTREAM_B200_SYNC=$m ./build/b200-stream --csv --gigabytes -s $s --only All -n 300 2>/dev/null | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,6 | tr '\n' ' ' | sed "s/^/s=$s b200 $m : /"; echo
      done
    done
  done
fi
if has smallncu; then
  # device-side durations at an L2-resident size: the reference's kernels vs ours
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/small_ref.csv ./oracle/_ref/cuda-stream -s 1048576 -n 6 --only All > /dev/null 2>&1
  BABELSTREAM_B200_SYNC=stream ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/small_b200.csv ./build/b200-stream -s 1048576 -n 6 --only All > /dev/null 2>&1
  BABELSTREAM_B200_SYNC=doorbell ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/small_b200_bell.csv ./build/b200-stream -s 1048576 -n 6 --only All > /dev/null 2>&1
  python - <<'PY'
import csv, collections
for f in ("small_ref", "small_b200", "small_b200_bell"):
    d = collections.defaultdict(list)
    try:
        rows = list(csv.reader(l for l in open(f"gpurun_out/{f}.csv") if l.startswith('"')))
    except Exception as e:
        print(f, "unreadable", e); continue
    hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value")
    for r in rows[1:]:
        d[r[ki][:60]].append(float(r[vi].replace(",", "")))
    print(f, {k: round(sorted(v)[len(v)//2]) for k, v in d.items()})
PY
fi
if has probe; then
  for n in 1024 1048576; do echo "== latency probe n=$n"; ./build/latency_probe $n; done 2>&1 | tee gpurun_out/latency_probe.txt
  BABELSTREAM_B200_HOSTPROF=1 ./build/b200-stream --csv -s 1024 -n 3000 2>&1 | grep -E "hostprof"
fi
if has managed; then
  for s in 33554432 536870912; do
    for m in device managed; do
      BABELSTREAM_B200_MEM=$m timeout 300 ./build/b200-stream --csv --gigabytes -s $s --only All -n 50 2>&1 | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,5 | tr '\n' ' ' | sed "s/^/s=$s mem=$m : /"; echo
    done
  done 2>&1 | tee gpurun_out/managed_vs_device.txt
fi
if has ncu29; then
  # north-star size: every kernel once, full metric set, through the unmodified driver (iteration 2 of 3)
  ./build/b200-stream -s 536870912 --only All -n 3 > gpurun_out/plain29.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k "regex:ew_kernel|dot_kernel" -s 7 -c 6 -f -o gpurun_out/prof_2p29 ./build/b200-stream -s 536870912 --only All -n 3 > gpurun_out/ncu29.log 2>&1
  echo "ncu29 rc=$?"; tail -2 gpurun_out/ncu29.log
  ncu -i gpurun_out/prof_2p29.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size > gpurun_out/ncu29_summary.csv 2>&1
  ncu --set full --clock-control none -k "regex:for_each_kernel|dot_kernel" -s 7 -c 6 -f -o gpurun_out/prof_2p29_ref ./oracle/_ref/cuda-stream -s 536870912 --only All -n 3 > gpurun_out/ncu29_ref.log 2>&1
  ncu -i gpurun_out/prof_2p29_ref.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size > gpurun_out/ncu29_ref_summary.csv 2>&1
fi
if has driver; then
  for s in 33554432 536870912; do
    for rep in 1 2; do
      timeout 200 ./oracle/_ref/cuda-stream --csv --gigabytes -s $s --only All > gpurun_out/drv_ref_cuda_${s}_$rep.csv 2>&1
      timeout 200 ./build/b200-stream --csv --gigabytes -s $s --only All > gpurun_out/drv_b200_${s}_$rep.csv 2>&1
    done
    BABELSTREAM_B200_SYNC=spin timeout 200 ./build/b200-stream --csv --gigabytes -s $s --only All > gpurun_out/drv_b200_spin_${s}.csv 2>&1
  done
  timeout 200 ./oracle/_ref/cuda-stream --csv --gigabytes -s 67108864 --float > gpurun_out/drv_ref_cuda_f32.csv 2>&1
  timeout 200 ./build/b200-stream --csv --gigabytes -s 67108864 --float > gpurun_out/drv_b200_f32.csv 2>&1
  grep -H -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," gpurun_out/drv_*.csv | cut -d, -f1,5,6
fi
echo "gpu_round done"
# (compute-sanitizer is closed on this pool: no memcheck/racecheck stages)
echo "gpu_round end"
