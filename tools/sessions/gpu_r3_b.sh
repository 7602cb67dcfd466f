#!/bin/bash
# round 3, call B: full GPU suite after the option-validation / doorbell-count changes, one default bench line
O=gpurun_out/r3b; mkdir -p $O
python -m pytest tests -x -q -m gpu -rs > $O/pytest_gpu.log 2>&1; echo "suite rc=$?" | tee -a $O/pytest_gpu.log
tail -n 25 $O/pytest_gpu.log
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
tail -c 1500 $O/bench_n1.json
