#!/bin/bash
# Short 8-GPU session (run under `gpurun --gpus 8`): the one-process-per-GPU path with the fused
# peer-memory dot (CUDA IPC mailboxes), weak scaling N = 2, 4, 8, the driver's config 5 with the default
# (p2p) combine, and mem=managed at 2^31.
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q -x -k "group or multi_gpu_managed or combine_modes or driver_multi" > gpurun_out/pytest_multi_8b.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_multi_8b.log
for G in 1 2 4 8; do
  if [ $G -eq 1 ]; then
    timeout 200 python bench.py --gpus 1 --steps 100 --warmup 5 --no-extras > gpurun_out/scale_${G}_fused.json 2> gpurun_out/scale_${G}_fused.err
  else
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $G --steps 100 --warmup 5 > gpurun_out/scale_${G}_fused.json 2> gpurun_out/scale_${G}_fused.err
  fi
  echo "bench N=$G rc=$?"; python -c "
import json
try:
    b=json.loads([l for l in open('gpurun_out/scale_${G}_fused.json') if l.startswith('{')][-1]); print('N=$G value',b['value'],'e2e',b['e2e']['value'],b['config'].get('dot_combine'),b['kernels_gbs'])
except Exception as e: print('parse failed',e); print(open('gpurun_out/scale_${G}_fused.err').read()[-1500:])
"
done
BABELSTREAM_B200_BENCH_COMBINE=nccl timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 100 --warmup 5 > gpurun_out/scale_8_nccl.json 2> gpurun_out/scale_8_nccl.err
python -c "
import json
b=json.loads([l for l in open('gpurun_out/scale_8_nccl.json') if l.startswith('{')][-1]); print('N=8 nccl value',b['value'],'e2e',b['e2e']['value'],b['kernels_gbs'])
"
for COMB in p2p nccl host; do
  BABELSTREAM_B200_NGPUS=8 BABELSTREAM_B200_DOT_COMBINE=$COMB timeout 100 ./build/b200-stream --csv --gigabytes -s $((33554432*8)) --only All 2>&1 | grep -E "^(Copy|Mul|Add|Triad|Dot|Nstream)," | cut -d, -f1,5,6 | tr '\n' ' ' | sed "s/^/G=8 $COMB /"; echo
done
BABELSTREAM_B200_NGPUS=8 timeout 200 ./build/b200-stream --gigabytes -s 2147483648 -n 20 --only All > gpurun_out/cfg5_8_p2p.log 2>&1; echo "cfg5 p2p rc=$?"; tail -8 gpurun_out/cfg5_8_p2p.log
( time BABELSTREAM_B200_NGPUS=8 BABELSTREAM_B200_MEM=managed timeout 200 ./build/b200-stream --gigabytes -s 2147483648 -n 20 --only All > gpurun_out/cfg5_8_managed.log 2>&1 ) 2>&1 | grep real; echo "cfg5 managed rc=$?"; tail -8 gpurun_out/cfg5_8_managed.log
echo "gpu_multi8 done"
