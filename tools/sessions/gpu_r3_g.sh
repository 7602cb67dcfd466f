#!/bin/bash
# round 3, call G (2 GPUs): final state -- whole GPU suite (worker-CPU registry test included), smoke(), one N=1 bench line
O=gpurun_out/r3g; mkdir -p $O
python -m pytest tests -x -q -m gpu -rs > $O/pytest_gpu_2gpu.log 2>&1; echo "suite rc=$?" | tee -a $O/pytest_gpu_2gpu.log
tail -n 6 $O/pytest_gpu_2gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 $O/smoke.log
python bench.py --no-extras > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"; cut -c1-400 $O/bench_n1.json
