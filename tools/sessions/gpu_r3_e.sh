#!/bin/bash
# round 3, call E: the default bench line with all extras (config-3 sweep included), timed
O=gpurun_out/r3e; mkdir -p $O
t0=$(date +%s)
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? in $(( $(date +%s) - t0 )) s"
python - <<'PY'
import json
l = json.loads(open("gpurun_out/r3e/bench_n1.json").read().strip().splitlines()[-1])
print("value", l["value"], "e2e", l["e2e"]["value"], "parity", l["parity"]["ok"])
sw = l["driver"].get("config3_size_sweep", {})
for r in sw.get("rows", []):
    print(r["log2n"], r["dtype"], r["regime"], r.get("b200_stream"), r.get("reference_cuda_stream"))
print(l["driver"].get("error"))
PY
