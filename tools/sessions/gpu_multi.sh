#!/bin/bash
# Multi-GPU session (run under `gpurun --gpus N`): sharded parity tests, torchrun bench, driver runs of
# BASELINE configs 4 and 5.   Usage: bash tools/gpu_multi.sh N [big]
set -u
N=${1:-2}
START_G=${START_G:-1}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/multi_smi_$N.txt
nvidia-smi topo -m >> gpurun_out/multi_smi_$N.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -k "multi_gpu or driver_multi" > gpurun_out/pytest_multi_$N.log 2>&1; echo "pytest multi rc=$?"; tail -5 gpurun_out/pytest_multi_$N.log
G=$START_G
while [ $G -le $N ]; do
  if [ $G -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 100 --warmup 5 --no-extras > gpurun_out/scale_$G.json 2> gpurun_out/scale_$G.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $G --steps 100 --warmup 5 > gpurun_out/scale_$G.json 2> gpurun_out/scale_$G.err
  fi
  echo "bench N=$G rc=$?"; python -c "
import json,sys
try:
    b=json.loads([l for l in open('gpurun_out/scale_$G.json') if l.startswith('{')][-1]); print('N=$G value',b['value'],'e2e',b['e2e']['value'],b['kernels_gbs'])
except Exception as e: print('parse failed',e); print(open('gpurun_out/scale_$G.err').read()[-1500:])
"
  G=$((G*2))
done
# BASELINE config 4: --only Triad / Nstream at 2^30 fp64, G = 1, 2, 4, 8 (strong scaling through the unmodified driver)
G=$START_G
while [ $G -le $N ]; do
  for K in Triad Nstream; do
    BABELSTREAM_B200_NGPUS=$G timeout 600 ./build/b200-stream --csv --gigabytes -s 1073741824 -n 30 --only $K 2> gpurun_out/cfg4_${K}_$G.err | grep -E "^(Triad|Nstream)," | sed "s/^/G=$G /"
  done
  G=$((G*2))
done
# all kernels, dot through NCCL vs host combine, at a per-GPU-friendly size
for COMB in nccl p2p host; do
  BABELSTREAM_B200_NGPUS=$N BABELSTREAM_B200_DOT_COMBINE=$COMB timeout 600 ./build/b200-stream --csv --gigabytes -s $((33554432*N)) --only All 2> gpurun_out/all_${COMB}_$N.err | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,5,6 | sed "s/^/G=$N $COMB /"
done
if [ "${2:-}" = "big" ]; then
  # BASELINE config 5: full suite, 2^31 fp64 over N GPUs (needs ~52 GB of host memory for validation)
  BABELSTREAM_B200_NGPUS=$N timeout 1500 ./build/b200-stream --gigabytes -s 2147483648 -n 20 --only All > gpurun_out/cfg5_$N.log 2>&1; echo "cfg5 rc=$?"; tail -12 gpurun_out/cfg5_$N.log
fi
echo "gpu_multi done"
