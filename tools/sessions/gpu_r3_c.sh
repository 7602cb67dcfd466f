#!/bin/bash
# round 3, call C: full GPU suite (option validation, no-alloc dot last), bench with the self-proving value region, reference arm
O=gpurun_out/r3c; mkdir -p $O
python -m pytest tests -x -q -m gpu -rs > $O/pytest_gpu.log 2>&1; echo "suite rc=$?" | tee -a $O/pytest_gpu.log
tail -n 12 $O/pytest_gpu.log
python bench.py --no-extras > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
l = json.loads(open("gpurun_out/r3c/bench_n1.json").read().strip().splitlines()[-1])
print("value", l["value"], "e2e", l["e2e"]["value"], "parity.ok", l["parity"]["ok"], "value_region", l["parity"]["value_region"])
print("config.dot_combine", l["config"]["dot_combine"])
PY
python bench.py --impl reference --steps 3 --warmup 3 --no-extras > $O/bench_ref.json 2>> $O/bench_n1.err; echo "ref rc=$?"; cut -c1-900 $O/bench_ref.json
