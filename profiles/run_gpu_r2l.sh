#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_gpu_r2l.log; cat gpurun_out/pytest_gpu_r2l.log
timeout 300 python profiles/e2e_probe.py > gpurun_out/e2e_probe_r2l.log 2>&1; cat gpurun_out/e2e_probe_r2l.log
CERES_LAUNCH_BANDS=1 timeout 300 python profiles/e2e_probe.py 0,4 >> gpurun_out/e2e_probe_r2l.log 2>&1; tail -2 gpurun_out/e2e_probe_r2l.log
CERES_TRACE_TIMING=1 timeout 100 python profiles/e2e_probe.py 0 2>&1 | grep -m3 "timing"
timeout 300 python profiles/kbench.py --share 1 --frames 20 2>&1 | tail -1
