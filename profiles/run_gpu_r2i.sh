#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_gpu_r2i.log; cat gpurun_out/pytest_gpu_r2i.log
timeout 300 python profiles/kbench.py --share 1 --frames 20 > gpurun_out/kbench_r2i.log 2>&1
timeout 300 python profiles/kbench.py --share 1 --frames 20 --workload dragon_1080p_shadow >> gpurun_out/kbench_r2i.log 2>&1
timeout 300 python profiles/kbench.py --share 0 --frames 5 --workload grid448_4k_shadow >> gpurun_out/kbench_r2i.log 2>&1
cat gpurun_out/kbench_r2i.log
