#!/bin/bash
# Round 2, multi-GPU session: bash tools/gpu_r2_multi.sh <G>   (run under `gpurun --gpus G`, G = 2 or 8)
# The @multi parity tests, bench.py at N = G (and N = 2 on an 8-GPU box), the in-process sharded driver.
set -u
G=${1:-2}
O=gpurun_out/r2m$G
mkdir -p $O
nvidia-smi -L | head -8
timeout 1500 python -m pytest tests -m gpu -q -rs > $O/pytest_gpu_${G}gpu.log 2>&1; echo "pytest rc=$?"; tail -8 $O/pytest_gpu_${G}gpu.log
run_bench() { # N extra-args
  local N=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 295$N bench.py --gpus $N --steps 100 --warmup 5 "$@" > $O/bench_$N.json 2> $O/bench_$N.err
  echo "bench N=$N rc=$?"
  python - <<PY
import json
try:
    b=json.loads([l for l in open('$O/bench_$N.json') if l.startswith('{')][-1])
    print('N=$N value',b['value'],'events',b['value_with_kernel_events'],'e2e',b['e2e']['value'],b['dot_combine'],'parity',b['parity']['ok'],b['parity']['dot_rel_err_max_over_ranks'])
    print(' kernels',b['kernels_gbs'])
    print(' strong',json.dumps(b.get('strong_scaling')))
except Exception as e:
    print('parse failed',e); print(open('$O/bench_$N.err').read()[-2500:])
PY
}
run_bench $G
if [ $G -eq 8 ]; then run_bench 2 --no-extras; run_bench 4 --no-extras; fi
echo "== in-process sharded driver, $G GPUs, 2^25 per GPU, --only All (GB/s)"
for rep in 1 2; do
  BABELSTREAM_B200_NGPUS=$G timeout 200 ./build/b200-stream --csv --gigabytes -s $((33554432*G)) --only All 2>&1 | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,5 | tr '\n' ' ' | tee -a $O/inprocess.txt; echo | tee -a $O/inprocess.txt
done
BABELSTREAM_B200_NGPUS=1 timeout 200 ./build/b200-stream --csv --gigabytes -s 33554432 --only All 2>&1 | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,5 | tr '\n' ' ' | sed 's/^/1 GPU: /' | tee -a $O/inprocess.txt; echo
BABELSTREAM_B200_NGPUS=$G BABELSTREAM_B200_LAUNCH=direct timeout 200 ./build/b200-stream --csv --gigabytes -s $((33554432*G)) --only All 2>&1 | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,5 | tr '\n' ' ' | sed 's/^/launch=direct: /' | tee -a $O/inprocess.txt; echo
for opt in "BABELSTREAM_B200_DOT_COMBINE=host" "BABELSTREAM_B200_DOT_COMBINE=nccl" "BABELSTREAM_B200_PIN=off"; do
  env $opt BABELSTREAM_B200_NGPUS=$G timeout 200 ./build/b200-stream --csv --gigabytes -s $((33554432*G)) --only All 2>&1 | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,5 | tr '\n' ' ' | sed "s/^/$opt: /" | tee -a $O/inprocess.txt; echo | tee -a $O/inprocess.txt
done
BABELSTREAM_B200_NGPUS=$G timeout 100 ./build/b200-stream -s $((33554432*G)) --only Triad -n 5 2>&1 | head -30 > $O/banner_${G}gpu.txt; grep -E "threads|combine" $O/banner_${G}gpu.txt
echo "gpu_r2_multi $G done"
