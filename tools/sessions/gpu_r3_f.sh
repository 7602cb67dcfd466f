#!/bin/bash
# round 3, call F (2 GPUs): caller-device handling. --device 1 through the unmodified driver must leave GPU 0 untouched
# (memory.used sampled while it runs); then the whole GPU suite.
O=gpurun_out/r3f; mkdir -p $O
nvidia-smi --query-gpu=index,memory.used --format=csv,noheader > $O/device1_mem.txt
( sleep 4; echo "-- during --device 1 run:"; nvidia-smi --query-gpu=index,memory.used --format=csv,noheader; nvidia-smi --query-compute-apps=gpu_bus_id,pid,used_memory --format=csv,noheader ) >> $O/device1_mem.txt &
./build/b200-stream --device 1 -s 134217728 -n 1500 > $O/device1_run.txt 2>&1; echo "device 1 run rc=$?"
wait
cat $O/device1_mem.txt; grep -E "Using|Triad|Dot" $O/device1_run.txt | head -5
python -m pytest tests -x -q -m gpu -rs > $O/pytest_gpu_2gpu.log 2>&1; echo "suite rc=$?" | tee -a $O/pytest_gpu_2gpu.log
tail -n 6 $O/pytest_gpu_2gpu.log
