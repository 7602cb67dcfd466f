#!/bin/bash
# Round 2, GPU session K (1 GPU): final verification of the tree as committed -- parity suite, smoke, both bench arms.
set -u
O=gpurun_out/r2k
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -rs > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
python bench.py --impl reference > $O/bench_reference.json 2> $O/bench_reference.err; echo "bench ref rc=$?"
python - <<PY
import json
b=json.loads(open('$O/bench_reference.json').read())
print('reference arm', b['value'], b['ms_per_step'], b['e2e_whole_run'])
PY
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"; tail -3 $O/bench_n1.err
python - <<PY
import json
b=json.loads(open('$O/bench_n1.json').read())
print('value',b['value'],'events',b['value_with_kernel_events'],'e2e',b['e2e']['value'],'parity',b['parity']['ok'])
print('kernels',b['kernels_gbs'])
print('north',json.dumps(b['roofline']['north_star']))
print('driver',json.dumps({k:v for k,v in b['driver'].items() if k!='north_star_triad_2^29'}))
print('whole',json.dumps(b['e2e_whole_run']))
print('clocks',b['clocks'])
PY
echo "gpu_r2_k done"
