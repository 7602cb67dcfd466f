#!/bin/bash
# round 3, call A: the no-allocation dot alone in a fresh process, then the whole GPU suite (it runs last there)
mkdir -p gpurun_out/r3a
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r3a/no_alloc_alone.log 2>&1; echo "alone rc=$?" | tee -a gpurun_out/r3a/no_alloc_alone.log
python -m pytest tests -x -q -m gpu > gpurun_out/r3a/pytest_gpu.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/r3a/pytest_gpu.log
tail -5 gpurun_out/r3a/no_alloc_alone.log gpurun_out/r3a/pytest_gpu.log
