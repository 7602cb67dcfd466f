#!/bin/bash
# round 3, call D (2 GPUs): the multi-GPU tests the 1-GPU box skips, bench at N=2 with extras (dot_combine through the driver)
O=gpurun_out/r3d; mkdir -p $O
python -m pytest tests -x -q -m gpu -rs > $O/pytest_gpu_2gpu.log 2>&1; echo "suite rc=$?" | tee -a $O/pytest_gpu_2gpu.log
tail -n 8 $O/pytest_gpu_2gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 50 --warmup 5 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench2 rc=$?"
python - <<'PY'
import json
l = json.loads(open("gpurun_out/r3d/bench_n2.json").read().strip().splitlines()[-1])
print("value", l["value"], "e2e", l["e2e"]["value"], "parity", l["parity"]["ok"], l["parity"]["value_region"]["ok"], "combine", l.get("dot_combine"))
ss = l.get("strong_scaling", {})
print(json.dumps(ss.get("driver_in_process", {}), indent=1)[:2500])
PY
tail -n 5 $O/bench_n2.err
